from .seed import set_seeds, rng_state, set_rng_state
from .logging import RoundLogger
from .checkpoint import save_checkpoint, load_checkpoint, latest_checkpoint
