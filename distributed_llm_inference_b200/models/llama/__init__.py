from .cache import PartialLlamaSinkCache  # noqa: F401
from .model import LlamaBlock  # noqa: F401
from .modules import OptimizedLlamaDecoderLayer, OptimizedLlamaInferenceAttention  # noqa: F401
