from .tinycnn import TinyCNN, flatten_state, unflatten_state, param_layout
from .vqc import VQC
