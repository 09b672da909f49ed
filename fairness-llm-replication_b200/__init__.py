"""B200-native scoring path of saakshipatel/fairness-llm-replication.

Import as ``fairness_llm_replication_b200``.  Sub-modules mirror the reference's module
names for the functions on the hot path:

* ``utils``                      -- calculate_individual_fairness, create_profile_pairs,
                                    calculate_demographic_parity, calculate_equal_opportunity,
                                    calculate_ndcg, calculate_exposure_ratio
* ``phase1_bias_detection``      -- measure_individual_fairness / _demographic_parity / _equal_opportunity,
                                    calculate_phase1_metrics, rescore_phase1_results
* ``phase2_cross_model_eval``    -- calculate_exposure_ratio, calculate_ndcg_per_group,
                                    calculate_pairwise_preference_ratio
* ``phase3_facter_mitigation``   -- generate_calibration_data, calculate_conformal_thresholds,
                                    apply_conformal_filtering, measure_quality_preservation
* ``phase3_final``               -- jaccard, measure_fairness
* ``ops``                        -- tensor-level entry points over the C ABI (include/flr_b200.h)
* ``pipeline``                   -- the sharded list-metric + conformal pipeline of the benchmark
"""
__version__ = "0.1.0"
