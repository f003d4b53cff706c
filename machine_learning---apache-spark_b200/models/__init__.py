"""Model families of the reference, same constructors and state_dict keys."""
from .cnn import FashionMNISTModel  # noqa: F401
from .lstm import LSTM  # noqa: F401
from .mlp import Multilayer_perceptor, mnist_mlp  # noqa: F401
from .transformer import Transformer  # noqa: F401
