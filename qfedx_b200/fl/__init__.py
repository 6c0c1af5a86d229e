from .accountant import RDPAccountant, rdp_sampled_gaussian, gaussian_epsilon_closed_form
from .dataset import FederatedDataset
from .sampler import sample_clients, inject_dropouts, participation_mask
from .server import FLConfig, FederatedRunner
from .trainers import TorchLocalTrainer, VQCFusedTrainer
from .api import federated_learning, quantum_federated_learning, prepare_quantum_features
