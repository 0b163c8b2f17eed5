"""Mirror of the reference's lib/data_utils for the hot path (bundles, .torch.idx/.bin reader); the rest
(fs, split, dataset_util, async_dataset, ...) resolves to the reference's own files when they are on sys.path."""
from .. import extend_package_path as _extend

_extend(__path__, "data_utils")
