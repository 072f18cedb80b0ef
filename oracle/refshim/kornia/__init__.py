"""Stand-in for kornia 0.2.2 (absent, cannot be installed offline); test infrastructure only.
Only the two functions the reference touches are provided (helpers/utils.py:97,101; model.py:199)."""
from . import filters  # noqa: F401
