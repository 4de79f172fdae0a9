"""CPU: the model API mirror (registry / factory / run / adapters / utils) and weight compatibility with the reference.

``tests/golden/model_*_state.json`` were written by tools/make_model_golden.py from the UNMODIFIED reference models
after ``torch.manual_seed(MODEL_SEED)``: equal names, shapes and sums prove that reference checkpoints load with
``strict=True`` and that the GPU parity tests share the reference's weights."""
import json
import os

import numpy as np
import pytest
import torch

import robustmvd_b200 as rb
from robustmvd_b200 import _lib, synthetic as syn, utils
from robustmvd_b200.models import helpers
from tools import cases

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_registry_lists_the_native_models():
    assert rb.list_models() == ["mvsnet_train", "robust_mvd", "robust_mvd_5M"]
    assert rb.list_models(trainable_only=True) == ["robust_mvd"]
    assert rb.has_model("robust_mvd") and not rb.has_model("nope")
    with pytest.raises(AssertionError):
        rb.create_model("nope")
    with pytest.raises(AssertionError):  # mvsnet.py:208: no pretrained weights exist
        rb.create_model("mvsnet_train", num_gpus=0)


@pytest.mark.parametrize("name", ["robust_mvd", "mvsnet_train"])
def test_seeded_state_dict_equals_the_reference(name):
    ref = json.load(open(os.path.join(GOLDEN, f"model_{name}_state.json")))
    torch.manual_seed(cases.MODEL_SEED)
    model = rb.create_model(name, pretrained=False, num_gpus=0)
    sd = model.state_dict()
    assert sorted(sd) == sorted(ref)
    for k, v in sd.items():
        assert list(v.shape) == ref[k]["shape"], k
        assert float(v.double().sum()) == pytest.approx(ref[k]["sum"], rel=1e-9, abs=1e-9), k
        assert float(v.double().abs().sum()) == pytest.approx(ref[k]["abssum"], rel=1e-9, abs=1e-9), k
    assert model.name == name and callable(model.run) and not model.training


def test_checkpoint_roundtrip_strict(tmp_path):
    torch.manual_seed(1)
    m = rb.create_model("robust_mvd", pretrained=False, num_gpus=0)
    path = str(tmp_path / "snapshot-iter-000000001.pt")
    torch.save({"model_state_dict": {"module." + k: v for k, v in m.state_dict().items()}}, path)
    m2 = rb.create_model("robust_mvd", weights=path, num_gpus=0)  # weights override pretrained, "module." stripped
    for (k, a), (_, b) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k


def test_robust_mvd_input_adapter():
    m = rb.create_model("robust_mvd", pretrained=False, num_gpus=0)
    s = syn.model_sample(100, 180, 2, seed=3, batch=1)
    out = m.input_adapter(**s)
    assert [tuple(i.shape) for i in out["images"]] == [(1, 3, 128, 192)] * 3      # next multiple of 64
    K = out["intrinsics"][0][0].numpy()
    assert np.allclose(K[0, 0], 0.58 * 180 * (192 / 180) / 192) and np.allclose(K[0, 2], 0.5)  # relative units
    assert np.allclose(K[1, 2], 0.5) and K[2, 2] == 1
    assert out["images"][0].dtype == torch.float32 and -0.41 < float(out["images"][0].min()) and float(out["images"][0].max()) < 0.61
    s2 = syn.model_sample(128, 192, 2, seed=3, batch=1)                            # no resize: exact arithmetic
    out2 = m.input_adapter(**s2)
    assert np.array_equal(out2["images"][1].numpy(), s2["images"][1] / 255.0 - 0.4)
    assert out2["keyview_idx"].tolist() == [0]


def test_mvsnet_input_adapter_and_schedule():
    torch.manual_seed(0)
    m = rb.create_model("mvsnet_train", pretrained=False, num_gpus=0)
    s = syn.model_sample(50, 90, 2, seed=4, batch=1, depth_range=(2.0, 40.0))
    out = m.input_adapter(**s)
    assert tuple(out["images"][0].shape) == (1, 3, 64, 96)                         # next multiple of 32
    d = m.depth_samples(out["depth_range"], 1, "cpu")
    assert tuple(d.shape) == (1, 256) and torch.equal(d[0], torch.linspace(2.0, 40.0, 256))
    assert torch.equal(m.depth_samples(None, 2, "cpu")[1], torch.linspace(0.2, 100, 256))
    poses = [p.clone() for p in out["poses"]]
    P = m.projection_matrices(out["poses"], out["intrinsics"], out["keyview_idx"])
    assert tuple(P.shape) == (1, 3, 4, 4)
    assert all(torch.equal(a, b) for a, b in zip(poses, out["poses"]))            # caller's poses are not modified
    K = out["intrinsics"][1][0] * torch.tensor([[0.25] * 3, [0.25] * 3, [1.0] * 3])
    assert torch.allclose(P[0, 1, :3, :4], K @ out["poses"][1][0, :3, :4])
    assert torch.allclose(P[0, 0] @ torch.cat([K @ out["poses"][0][0, :3, :4], torch.tensor([[0, 0, 0, 1.0]])]),
                          torch.eye(4), atol=1e-5)


def test_forward_has_no_cpu_fallback():
    m = rb.create_model("robust_mvd", pretrained=False, num_gpus=0)
    s = syn.model_sample(64, 64, 1, seed=5)
    with pytest.raises(_lib.RmvdError):
        m.run(**s)


def test_batch_helpers_and_index_utils():
    s = syn.model_sample(8, 8, 2, seed=6)
    images, key, poses, K, rng = helpers.add_batch_dim(s["images"], s["keyview_idx"], s["poses"], s["intrinsics"])
    assert images[0].shape == (1, 3, 8, 8) and key.tolist() == [0] and rng is None
    out = helpers.remove_batch_dim(({"depth": np.zeros((1, 1, 4, 4))}, {"all": [np.ones((1, 2))], "none": None}))
    assert out[0]["depth"].shape == (1, 4, 4) and out[1]["all"][0].shape == (2,) and out[1]["none"] is None
    views = [torch.full((2, 1), float(v)) for v in range(3)]
    assert utils.select_by_index(views, 1) is views[1]
    assert utils.select_by_index(views, torch.tensor([2, 0])).flatten().tolist() == [2.0, 0.0]
    ex = utils.exclude_index(views, torch.tensor([2, 0]))
    assert [e.flatten().tolist() for e in ex] == [[0.0, 1.0], [1.0, 2.0]]
    assert len(utils.exclude_index(views, 0)) == 2
    assert utils.compute_depth_range() == (0.1, 100.0)                             # project-wide default range
    assert utils.compute_depth_range(depth=np.array([0.0, 2.0, 5.0])) == (2.0, 5.0)
    t = utils.to_torch(({"a": np.ones(2, np.float32)}, [np.int64(3)], None, "s"))
    assert isinstance(t, list) and isinstance(t[0]["a"], torch.Tensor) and t[2] is None and t[3] == "s"
    assert isinstance(utils.to_numpy(t)[0]["a"], np.ndarray)


def test_prepare_custom_model():
    class Tiny(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.p = torch.nn.Parameter(torch.ones(1))

        def input_adapter(self, images, **_):
            return {"images": utils.to_torch(images)}

        def forward(self, images):
            return {"depth": images[0][:, :1] * self.p}, {}

        def output_adapter(self, out):
            return utils.to_numpy(out[0]), utils.to_numpy(out[1])

    m = rb.prepare_custom_model(Tiny(), num_gpus=0)
    pred, aux = m.run(images=[np.full((3, 4, 4), 2.0, np.float32)], keyview_idx=0)
    assert pred["depth"].shape == (1, 4, 4) and float(pred["depth"].mean()) == 2.0 and aux == {}
