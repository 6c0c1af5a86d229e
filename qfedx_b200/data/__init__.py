from .idx import read_idx_images, read_idx_labels, write_idx_images, write_idx_labels
from .partition import iid_indices, dirichlet_indices, partition, partition_summary, take
from .features import (downsample_images, pool_features, angle_features, amplitude_normalize,
                       PCAReducer, images_to_angles, images_to_amplitudes)
from .synthetic import synthetic_mnist, synthetic_uint8_corpus, synthetic_images_for_labels
from .preprocess import preprocess_mnist, load_processed, Prepared, stratified_split
from .iris import load_iris, prepare_iris, iris_angles
