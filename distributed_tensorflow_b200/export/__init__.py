from .saved_model import SavedModelBuilder, export_saved_model, load_saved_model, PREDICT_METHOD_NAME, \
    DEFAULT_SERVING_SIGNATURE_DEF_KEY, SERVING  # noqa: F401
from .session_bundle import export_session_bundle  # noqa: F401
