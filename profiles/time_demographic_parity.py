import sys, torch
sys.path.insert(0, ".")
from fairness_llm_replication_b200 import ops, pipeline
U, V, K, vocab = 1250000, 31, 50, 3706
_, _, lists, lens = ops.synth_lists(pipeline.SEED, 0, U, V, K, vocab, "cuda:0")
for _ in range(2): h = ops.group_item_histogram(lists, lens, vocab=vocab, n_groups=V)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); h = ops.group_item_histogram(lists, lens, vocab=vocab, n_groups=V); b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b); gb = lists.numel()*4/1e9
print("histogram 1.25M users: %.2f ms, %.0f GB/s of list reads" % (ms, gb/ms*1e3))
a.record(); d, v = ops.js_distance_matrix(h); b.record(); torch.cuda.synchronize()
print("js matrix (465 pairs): %.3f ms; score %.6f" % (a.elapsed_time(b), 1 - float(d[v.bool()].mean())))
# the same histogram through the per-group kernel (explicit group ids) must agree exactly
g = (torch.arange(U * V, device="cuda:0") % V).to(torch.uint8)
h2 = ops.group_item_histogram(lists.view(U * V, K), lens.view(-1), vocab=vocab, n_groups=V, groups=g)
torch.cuda.synchronize()
print("dense == per-group kernel:", bool(torch.equal(h, h2)), int(h.sum()), int(lens.clamp(max=K).sum()))
