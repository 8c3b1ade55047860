"""sgc_b200 -- B200-native (sm_100a) implementation of SGC's feature-propagation hot path:
``normalization.aug_normalized_adjacency`` and ``utils.sgc_precompute`` of bellaj09/SGC, behind the
reference's own Python API.  CUDA kernels live in sgc_b200/csrc (C ABI: include/sgc_b200.h);
PyTorch is used for device memory, streams and torch.distributed only."""
from .graph import (SgcGraph, augment_csr, finish_normalization, from_scipy, from_torch_sparse,  # noqa: F401
                    normalize_csr)
from .features import StandardizedFeatures, col_stats, standardize_, to_padded  # noqa: F401
from . import propagate  # noqa: F401  (module: propagate.propagate() is the stream-ordered K-hop call)
from .propagate import hop, sgc_precompute  # noqa: F401
from . import pipeline  # noqa: F401
from .pipeline import precompute_host  # noqa: F401
from .normalization import aug_normalized_adjacency, fetch_normalization, row_normalize  # noqa: F401

__version__ = "0.1.0"
