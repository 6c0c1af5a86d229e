import math
import os
import random

import numpy as np
import pytest
import torch

from qfedx_b200 import data as D
from qfedx_b200.oracle import encoders as enc

REF_RAW = "/root/reference/dataset/raw"


def test_idx_roundtrip_and_magic_check(tmp_path):
    imgs = np.random.default_rng(0).integers(0, 256, (7, 28, 28), dtype=np.uint8)
    labs = np.arange(7, dtype=np.uint8)
    D.write_idx_images(tmp_path / "i.idx3", imgs)
    D.write_idx_labels(tmp_path / "l.idx1", labs)
    np.testing.assert_array_equal(D.read_idx_images(tmp_path / "i.idx3"), imgs)
    np.testing.assert_array_equal(D.read_idx_labels(tmp_path / "l.idx1"), labs)
    with pytest.raises(ValueError):
        D.read_idx_images(tmp_path / "l.idx1")


def test_iid_partition_sizes_and_remainder():
    random.seed(42)
    parts = D.iid_indices(16760, 4)
    assert [len(p) for p in parts] == [4190] * 4
    random.seed(1)
    parts = D.iid_indices(103, 4)
    assert [len(p) for p in parts] == [25, 25, 25, 28]          # last client takes the tail
    assert sorted(np.concatenate(parts).tolist()) == list(range(103))


def test_iid_matches_reference_shuffle_rule():
    random.seed(42)
    ours = D.iid_indices(50, 3)
    random.seed(42)
    order = list(range(50)); random.shuffle(order)
    assert ours[0].tolist() == order[:16] and ours[2].tolist() == order[32:]


def test_dirichlet_partition_properties():
    random.seed(0); np.random.seed(0)
    y = np.repeat(np.arange(3), [500, 700, 300])
    parts = D.dirichlet_indices(y, 5, alpha=0.5)
    allidx = np.concatenate(parts)
    assert sorted(allidx.tolist()) == list(range(1500))         # disjoint cover
    for p in parts:                                               # class-sorted order within a client
        assert np.all(np.diff(y[p]) >= 0)


def test_dirichlet_works_for_non_contiguous_labels():
    random.seed(0); np.random.seed(0)
    y = np.repeat(np.array([3, 5, 8]), 100)
    parts = D.dirichlet_indices(y, 4, alpha=1.0)
    assert sum(len(p) for p in parts) == 300                     # the reference would give 0 here


def test_partition_keeps_torch_tensors():
    X, y = D.synthetic_mnist(64, 3)
    random.seed(0)
    cl = D.partition(X, y, 4, "iid")
    assert all(isinstance(c[0], torch.Tensor) and c[1].dtype == torch.int64 for c in cl)


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_RAW, "train-labels.idx1-ubyte")), reason="reference labels absent")
def test_preprocess_reproduces_reference_shapes(tmp_path):
    random.seed(42); np.random.seed(42)
    prep = D.preprocess_mnist(REF_RAW, str(tmp_path / "proc"), synthetic_fallback=True, verbose=False)
    assert prep.train[0].shape == (16760, 1, 28, 28) and prep.val[0].shape[0] == 1863 and prep.test[0].shape[0] == 3147
    assert [len(c[0]) for c in prep.clients] == [4190] * 4
    assert prep.train[0].dtype == torch.float32 and prep.train[1].dtype == torch.int64
    assert 0.0 <= float(prep.train[0].min()) and float(prep.train[0].max()) <= 1.0
    back = D.load_processed(str(tmp_path / "proc"))
    assert torch.equal(back["test"][1], prep.test[1])


def test_preprocess_without_files_returns_none(tmp_path):
    assert D.preprocess_mnist(str(tmp_path), str(tmp_path / "p"), verbose=False) is None


def test_preprocess_synthetic_small(tmp_path):
    random.seed(0)
    prep = D.preprocess_mnist(str(tmp_path), str(tmp_path / "p"), digits=(3, 5), num_clients=3, remap_labels=True,
                              synthetic_fallback=True, synthetic_scale=0.01, verbose=False)
    assert set(prep.train[1].unique().tolist()) == {0, 1}
    assert len(prep.clients) == 3


def test_torch_features_match_oracle():
    g = torch.Generator().manual_seed(0)
    X = torch.rand(5, 1, 28, 28, generator=g, dtype=torch.float64)
    small = D.downsample_images(X[:, 0], (4, 4))
    for b in range(5):
        np.testing.assert_allclose(small[b].numpy(), enc.downsample_image(X[b, 0].numpy()), atol=1e-12)
    flat = small.reshape(5, 16)
    np.testing.assert_allclose(D.pool_features(flat, 4)[2].numpy(), enc.pool_to_n_features(flat[2].numpy(), 4), atol=1e-12)
    np.testing.assert_allclose(D.pool_features(flat, 5)[1].numpy(), enc.pool_to_n_features(flat[1].numpy(), 5), atol=1e-12)
    np.testing.assert_allclose(D.angle_features(flat, 4)[3].numpy(), enc.angle_features(flat[3].numpy(), 4), atol=1e-12)
    np.testing.assert_allclose(D.images_to_angles(X, 4)[0].numpy(), enc.angle_features(flat[0].numpy(), 4), atol=1e-12)
    amp = D.amplitude_normalize(flat)
    np.testing.assert_allclose(amp[4].numpy(), enc.normalize_for_amplitude(flat[4].numpy()), atol=1e-12)
    z = D.amplitude_normalize(torch.zeros(2, 8))
    assert torch.allclose(z, torch.full((2, 8), 1 / math.sqrt(8)))
    assert torch.all(D.angle_features(torch.ones(2, 4)) == 0)


def test_pca_reducer_roundtrip():
    X = torch.randn(200, 1, 6, 6, generator=torch.Generator().manual_seed(0))
    p = D.PCAReducer(4).fit(X)
    Z = p.transform(X)
    assert Z.shape == (200, 4)
    var = Z.var(0)
    assert torch.all(var[:-1] >= var[1:] - 1e-4)                 # components ordered by variance
    q = D.PCAReducer.from_state_dict(p.state_dict())
    assert torch.allclose(q.transform(X), Z)


def test_plots_write_files_or_decline_without_matplotlib(tmp_path):
    """R: src/CFed/Preprocess.py:71-134 - file output only; a missing matplotlib turns the calls into no-ops."""
    import importlib.util
    from qfedx_b200.data import partition, synthetic_mnist
    from qfedx_b200.data.plots import class_distribution_table, plot_class_distribution, visualize_client_data
    X, y = synthetic_mnist(120, 3, seed=0)
    clients = partition(X, y, 4, "iid")
    table = class_distribution_table(clients)
    assert table.count("\n") >= 4                                   # header + one line per client
    have = importlib.util.find_spec("matplotlib") is not None
    a = visualize_client_data(clients, str(tmp_path / "samples.png"))
    b = plot_class_distribution(clients, str(tmp_path / "dist.png"))
    assert a == have and b == have
    assert (tmp_path / "dist.png").exists() == have
