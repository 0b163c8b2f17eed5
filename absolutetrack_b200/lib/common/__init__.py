"""Mirror of the reference's lib/common for the hot path (cameras, crops, hand model, FK/skinning, metrics)."""
from .. import extend_package_path as _extend

_extend(__path__, "common")
