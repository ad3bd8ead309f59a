"""`knn` under the reference's import path (gs_init_compare/utils/runner_utils.py:143-147)."""
from ..knn import knn  # noqa: F401
