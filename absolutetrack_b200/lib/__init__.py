"""Host-side mirror of the reference's `lib` package for the per-frame inference hot path.

Importable as `absolutetrack_b200.lib` or, with absolutetrack_b200/ on sys.path, as `lib` (the name
the reference's entry scripts use).  Only the modules of the hot path are mirrored.  Everything
else the entry scripts import (`lib.data_utils.fs`, `.split`, `.dataset_util`, `.async_dataset`,
... : host plumbing) falls through to the reference's own copy: when another `lib/` directory is
found on sys.path (the reference root appended AFTER this package's directory), its matching
sub-directories are appended to the `__path__` of this package and of every sub-package, so
`import lib.data_utils.fs` finds the reference file while `import lib.tracker.tracker` finds the
mirror.  Nothing is imported from there by this package itself.
"""
import os
import sys

try:
    from .. import _cabi  # imported as absolutetrack_b200.lib
except ImportError:  # imported as top-level `lib`
    _root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    if _root not in sys.path:
        sys.path.append(_root)
    from absolutetrack_b200 import _cabi  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))


def reference_lib_dirs():
    """`lib/` directories of a reference checkout on sys.path (not this mirror), in sys.path order."""
    found = []
    for entry in sys.path:
        cand = os.path.abspath(os.path.join(entry or os.getcwd(), "lib"))
        if cand != _HERE and cand not in found and os.path.isdir(os.path.join(cand, "data_utils")):
            found.append(cand)
    return found


def extend_package_path(path, *sub):
    """Append the reference's copy of sub-package `sub` to the package search path `path`."""
    for ref in reference_lib_dirs():
        cand = os.path.join(ref, *sub)
        if os.path.isdir(cand) and cand not in path:
            path.append(cand)
    return path


def refresh_reference_paths() -> None:
    """Re-scan sys.path (call it after adding the reference root to sys.path at run time)."""
    me = sys.modules[__name__]
    extend_package_path(me.__path__)
    for name in ("batched_dataset", "common", "data_utils", "models", "tracker"):
        mod = sys.modules.get(f"{__name__}.{name}")
        if mod is not None:
            extend_package_path(mod.__path__, name)


extend_package_path(__path__)
