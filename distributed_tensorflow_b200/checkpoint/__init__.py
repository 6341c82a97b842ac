from .bundle import BundleReader, BundleWriter, merge_bundles, DTYPE_TO_TF, TF_TO_DTYPE  # noqa: F401
from .saver import Saver, latest_checkpoint, get_checkpoint_state, update_checkpoint_state, checkpoint_exists  # noqa: F401
