"""ytk-learn_b200: the B200-native GBDT tree maker behind ytk-learn's ITreeMaker seam.

Host-side mirror (Python, for tests and benches) of the reference interface on top of
the C ABI in include/ytk_gbdt.h; the product is libytkgbdt.so (csrc/).  The directory
name is not an importable identifier; load it with `_pkg.load()` from the repo root,
which registers it as module `ytk_learn_b200`.
"""
from .binding import (LIB_PATH, YtkError, Params, Split, TimeStats, Tree, avg_assign,  # noqa: F401
                      lib, merge_splits, nccl_unique_id, normalize_max_leaf_cnt)
from .tree_maker import GpuTreeMaker, make_params  # noqa: F401
