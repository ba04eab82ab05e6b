"""melissa_b200: B200-native variational-Bayes hot path of Melissa.

Public surface mirrors the reference's R functions for this path:
`melissa` / `melissa_vb` (R/melissa_vb.R:86), `create_rbf_object`
(BPRMeth), `impute_test_met`, `eval_imputation_performance`,
`eval_cluster_performance` (R/utils.R).  The arithmetic runs in
libmelissa_b200.so (CUDA, sm_100a); there is no CPU fallback.
"""
from . import _lib
from .data import MelissaData, create_rbf_object, from_dict, from_nested, to_nested
from .engine import Engine, MultiEngine, progress, set_option
from .evaluate import (cluster_ari, cluster_error, eval_cluster_performance, eval_imputation_performance,
                       impute_test_met)
from .filters import filter_by_coverage_across_cells, filter_by_cpg_coverage, filter_by_variability
from .vb import MelissaResult, finalize, melissa, melissa_inner, melissa_vb, shard_bounds, vb_loop

__all__ = ["MelissaData", "create_rbf_object", "from_dict", "from_nested", "to_nested", "Engine", "MultiEngine", "progress", "set_option",
           "cluster_ari", "cluster_error", "eval_cluster_performance", "eval_imputation_performance", "impute_test_met",
           "filter_by_coverage_across_cells", "filter_by_cpg_coverage", "filter_by_variability", "MelissaResult", "finalize", "melissa", "melissa_inner", "melissa_vb", "shard_bounds", "vb_loop"]
