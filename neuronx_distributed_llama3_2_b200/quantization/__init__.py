"""int8 weight-only quantization (reference ``src/neuronx_distributed/quantization/*.py``)."""
from .dequantize import direct_cast_dequantize, scale_dequantize  # noqa: F401
from .observer import PerChannelAbsMaxObserver  # noqa: F401
from .quantization_config import (BASE_QCONFIG_DICT_TYPE, QuantizationType, QuantizedDtype,  # noqa: F401
                                  get_default_custom_qconfig_dict, get_default_per_channel_custom_qconfig_dict)
from .quantization_layers import (BaseQuantizeParallelLinear, QuantizedColumnParallel,  # noqa: F401
                                  QuantizedExpertFusedColumnParallel, QuantizedExpertFusedRowParallel,
                                  QuantizedParallelLinearLayerStateDictAdaptor, QuantizedRowParallel)
from .quantization_utils import (convert_qint8_to_int8_state_dict, extract_q_scale,  # noqa: F401
                                 extract_q_scale_per_channel, extract_q_scale_per_tensor,
                                 quantize_per_channel_symmetric, quantize_per_tensor_symmetric,
                                 quantize_pytorch_model_per_channel_symmetric,
                                 quantize_pytorch_model_per_tensor_symmetric, quantize_state_dict)
from .quantize import convert  # noqa: F401
