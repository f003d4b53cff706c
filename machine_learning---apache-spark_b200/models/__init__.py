"""Model families of the reference, same constructors and state_dict keys."""
from .transformer import Transformer  # noqa: F401
