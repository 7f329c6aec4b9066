from .modules import MLP, Head, MultiHeadAttention, Block  # noqa: F401
from .transformer import Transformer  # noqa: F401
from .layout import ParamLayout  # noqa: F401
