"""On-disk formats: gsb200.io against outputs of the reference's own loaders on the
same files (tests/golden/io_sequence.npz, written by tests/golden/make_golden_io.py from shared.py /
densify.py run unmodified).  CPU only: the loaders take a device argument."""
import json
import os

import numpy as np
import pytest
import torch

from gsb200 import io as gio

GOLD = os.path.join(os.path.dirname(__file__), "golden", "io_sequence.npz")


@pytest.fixture(scope="module")
def sequence(tmp_path_factory):
    g = np.load(GOLD)
    root = tmp_path_factory.mktemp("seq")
    for key in g.files:
        if key.startswith("file:"):
            path = root / key[5:]
            path.parent.mkdir(parents=True, exist_ok=True)
            path.write_bytes(g[key].tobytes())
    return root, g


def test_views_match_reference_loader(sequence):
    root, g = sequence
    meta = gio.load_dataset_metadata(root)
    assert set(meta) >= {"w", "h", "k", "w2c", "fn"}
    for t in range(len(meta["fn"])):
        views = gio.load_timestep_views(meta, t, root, device="cpu")
        assert [v.camera_index for v in views] == list(range(len(meta["fn"][t])))
        for v in views:
            s, p = v.render_settings, f"t{t}_c{v.camera_index}_"
            # bit-exact: same float32 operations in the same order
            assert np.array_equal(s.viewmatrix.numpy(), g[p + "viewmatrix"])
            assert np.array_equal(s.projmatrix.numpy(), g[p + "projmatrix"])
            assert np.array_equal(s.campos.numpy(), g[p + "campos"])
            assert np.array_equal(s.bg.numpy(), g[p + "bg"])
            sc = g[p + "scalars"]
            assert [s.image_height, s.image_width, s.tanfovx, s.tanfovy, s.scale_modifier, s.sh_degree,
                    float(s.prefiltered)] == list(sc)
            assert np.array_equal(v.image.numpy(), g[p + "image"])
            assert np.array_equal(v.segmentation_mask.numpy(), g[p + "segmentation_mask"])
            assert v.image.shape == (3, meta["h"], meta["w"]) and v.image.dtype == torch.float32


def test_initialize_parameters_matches_reference(sequence):
    root, g = sequence
    params = gio.initialize_parameters(root, device="cpu")
    assert tuple(params) == gio.PARAMETER_KEYS   # same 8 entries in the same order (densify.py:29-57)
    for k, v in params.items():
        assert isinstance(v, torch.nn.Parameter) and v.requires_grad and v.dtype == torch.float32
        assert np.array_equal(v.detach().numpy(), g["param_" + k]), k
    meta = gio.load_dataset_metadata(root)
    assert gio.get_scene_radius(meta) == float(g["scene_radius"])


def test_parameter_file_round_trip(sequence, tmp_path):
    root, _ = sequence
    params = gio.initialize_parameters(root, device="cpu")
    seq = tmp_path / "data" / "basketball"
    seq.mkdir(parents=True)
    path = gio.export_parameters(seq, params)
    assert path.name == "densified_initial_gaussian_cloud_parameters.pth"
    # the file is what the reference's train.py:155-163 reads: a plain torch.load of a dict of nn.Parameter
    raw = torch.load(path, weights_only=False)
    assert all(isinstance(v, torch.nn.Parameter) for v in raw.values())
    back = gio.load_densified_initial_parameters(tmp_path / "data", "basketball", device="cpu")
    for k in gio.PARAMETER_KEYS:
        assert not back[k].requires_grad
        assert torch.equal(back[k].data, params[k].data)
    torch.save({"means": params["means"]}, path)
    with pytest.raises(KeyError):
        gio.load_densified_initial_parameters(tmp_path / "data", "basketball", device="cpu")


def test_stack_views_layout(sequence):
    root, _ = sequence
    meta = gio.load_dataset_metadata(root)
    views = gio.load_timestep_views(meta, 0, root, device="cpu")
    cams, images, masks = gio.stack_views(views)
    n = len(views)
    assert cams["viewmatrix"].shape == (n, 4, 4) and cams["projmatrix"].shape == (n, 4, 4)
    assert len(cams["tanfovx"]) == n and cams["image_width"] == meta["w"] and cams["image_height"] == meta["h"]
    assert images.shape == (n, 3, meta["h"], meta["w"]) and masks.shape == images.shape
    assert torch.equal(images[1], views[1].image)
    # fg + bg channels of the mask partition the image (shared.py:163-169)
    assert torch.equal(masks[:, 0] + masks[:, 2], torch.ones_like(masks[:, 0])) and not masks[:, 1].any()


def test_write_sequence_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    meta = {"w": 8, "h": 6, "k": [[[[5.0, 0, 4], [0, 5.0, 3], [0, 0, 1]]]], "w2c": [[np.eye(4).tolist()]],
            "fn": [["0/000000.jpg"]]}
    mask = (rng.uniform(size=(6, 8)) > 0.5).astype(np.uint8)
    cloud = rng.uniform(size=(10, 7))
    gio.write_sequence(tmp_path, meta, [[rng.integers(0, 256, size=(6, 8, 3), dtype=np.uint8)]], [[mask]], cloud)
    assert json.load(open(tmp_path / "train_meta.json"))["fn"] == meta["fn"]
    v = gio.load_timestep_views(gio.load_dataset_metadata(tmp_path), 0, tmp_path, device="cpu")[0]
    assert np.array_equal(v.segmentation_mask[0].numpy(), mask.astype(np.float32))   # PNG is lossless
    assert np.array_equal(np.load(tmp_path / "init_pt_cld.npz")["data"], cloud)
