"""Score networks on the B200 engine.  Importing registers 'ncsnpp' and 'ddpm' (as the reference's
models/__init__ + run_lib imports do)."""
from . import utils  # noqa: F401
from . import ncsnpp, ddpm  # noqa: F401
