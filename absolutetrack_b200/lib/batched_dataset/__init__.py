"""GPU side of the reference's lib/batched_dataset (config 3, torch_data sequences)."""
from .. import extend_package_path as _extend

_extend(__path__, "batched_dataset")
