"""Drop-in for the reference's top-level ``utils`` module (/root/reference/utils.py): the names
``citation.py:7``, ``reddit.py:7`` and ``tuning.py:8`` import resolve here to the GPU
implementation.  See sgc_b200/utils.py."""
from sgc_b200.utils import (  # noqa: F401
    load_citation, load_reddit_data, loadRedditFromNPZ, parse_index_file, preprocess_citation,
    set_seed, sgc_precompute, sparse_mx_to_torch_sparse_tensor,
)
from sgc_b200.normalization import fetch_normalization, row_normalize  # noqa: F401  (utils.py:7)
