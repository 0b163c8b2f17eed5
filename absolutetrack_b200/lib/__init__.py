"""Host-side mirror of the reference's `lib` package for the per-frame inference hot path.

Importable as `absolutetrack_b200.lib` or, with absolutetrack_b200/ on sys.path, as `lib`
(the name the reference's entry scripts use).
"""
try:
    from .. import _cabi  # imported as absolutetrack_b200.lib
except ImportError:  # imported as top-level `lib`
    import os
    import sys

    _root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    if _root not in sys.path:
        sys.path.append(_root)
    from absolutetrack_b200 import _cabi  # noqa: F401
