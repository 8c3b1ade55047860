"""Drop-in for the reference's top-level ``normalization`` module (/root/reference/normalization.py):
put this repository ahead of the reference on ``sys.path`` and ``citation.py`` / ``reddit.py`` /
``tuning.py`` pick up the GPU implementation unchanged.  See sgc_b200/normalization.py."""
from sgc_b200.normalization import aug_normalized_adjacency, fetch_normalization, row_normalize  # noqa: F401
