"""Multi-GPU parity on real hardware: N ranks over NCCL score their shards, rank 0 re-scores the union on
its own GPU, and every global quantity must agree (pipeline.multi_gpu_parity).  Needs at least two GPUs:
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`; skipped on a one-GPU box.  The same
check runs inside `bench.py` under torchrun (its `check` object)."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, users_per_rank, list_len, out):
    import torch.distributed as dist
    from fairness_llm_replication_b200 import pipeline
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        verdict = pipeline.multi_gpu_parity(pipeline.PipelineConfig(list_len=list_len), rank, world, dev,
                                            users_per_rank=users_per_rank)
        if rank == 0:
            out.put(verdict)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("users_per_rank,list_len", [(4099, 50), (1000, 20)])
def test_sharded_run_equals_the_union_on_one_gpu(users_per_rank, list_len):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, users_per_rank, list_len, out)) for r in range(world)]
    for p in procs:
        p.start()
    verdict = out.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert verdict["ranks"] == world
    assert verdict["multi_gpu_parity"], verdict["checks"]
