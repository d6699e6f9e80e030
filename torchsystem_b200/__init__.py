"""torchsystem_b200: a B200-native (sm_100a) engine behind TorchSystem's train / validate / iterate handlers.

Importing this package loads libtsb200.so; there is no CPU, PyTorch-eager or Triton fallback."""
from ._lib import lib as _lib_handle, TsError  # noqa: F401  (raises ImportError if the library is not built)
from .classifier import B200Classifier
from .metrics import B200Metrics
from .loader import DeviceLoader, PinnedLoader
from .lowering import lower_classifier
from .modules import MLP
from . import training

__all__ = ['B200Classifier', 'B200Metrics', 'DeviceLoader', 'PinnedLoader', 'lower_classifier', 'MLP', 'TsError', 'training']
