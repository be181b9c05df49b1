from .tinygpt import TinyGPT, TransformerBlock, SelfAttention  # noqa: F401
from .layout import ParamLayout, ParamSpec, build_layout  # noqa: F401
