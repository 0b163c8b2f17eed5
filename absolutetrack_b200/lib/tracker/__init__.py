"""Mirror of the reference's lib/tracker API over libabsolutetrack_b200.so."""
from .. import extend_package_path as _extend

_extend(__path__, "tracker")
