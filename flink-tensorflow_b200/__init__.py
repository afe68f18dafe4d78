"""flink-tensorflow_b200: B200-native runner for flink-tensorflow's per-record model-invocation path.

Importing this package loads libftf_b200.so (the CUDA runner behind the C ABI of include/ftf.h) and fails
loudly when it has not been built; there is no CPU or pure-Python path.
"""
from . import _abi

_abi.load_library()

from .host import (DataType, GenericModel, GraphMethod, HalfPlusTwo, ImageNormalization,  # noqa: E402,F401
                   ImageNormalizationMethod, InceptionModel, LabeledImage, LabelMethod, MiniCluster,
                   ModelFunction, PredictMethod, RegressionMethod, RichModel, Runner, TensorFlowModel,
                   TensorName, TensorValue, VersionMismatchException, bytes_to_tensor, default_config, example,
                   map_with_model, messages_to_tensor)

__all__ = [n for n in dir() if not n.startswith("_")]
