"""B200-native drop-in for the training step of cifar10cnn.py.

Importing the package does not need a GPU (the CLI parser, role logic and the ABI loader are
CPU code); every compute entry point does, and raises without one."""
from . import _lib  # noqa: F401
from .engine import PARAM_NAMES, PARAM_SCOPES, TrainStepEngine  # noqa: F401

# (mode, tolerance) pairs exercised by __graft_entry__.smoke(); tolerance is max-abs-normalised
# relative error against the fp64 oracle after one SGD step.
SMOKE_MODES = (("fp32", 2e-5), ("bf16", 2e-2))
from .exchange import GradientExchange, P2PExchange  # noqa: E402,F401
# mode bench.py and the CLI use unless told otherwise
DEFAULT_MODE = "bf16"
from . import checkpoint, data, roles, session  # noqa: E402,F401
