"""Mirror of the reference's lib/models API over libabsolutetrack_b200.so."""
from .. import extend_package_path as _extend

_extend(__path__, "models")
