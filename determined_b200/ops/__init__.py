"""sm_100a kernel library bindings and the nn.Module / autograd layer built on them."""
from . import _native, functional  # noqa: F401
