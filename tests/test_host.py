"""CPU tests of the host layer: ABI surface, TorchScript lowering, scalar arithmetic, file contract,
multi-rank sharding/gather (gloo, world_size 2).  No kernel is launched here."""

import os
import re
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol():
    from wsinsight_b200 import _cabi

    header = (ROOT / "include" / "wsi_b200.h").read_text()
    declared = set(re.findall(r"\b(wsi_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    lib = _cabi.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in wsi_b200.h but not exported"
    assert declared == set(_cabi.EXPORTED_SYMBOLS), declared ^ set(_cabi.EXPORTED_SYMBOLS)
    assert b"sm_100a" in lib.wsi_version()


def test_library_contains_blackwell_sass():
    """The shipped .so must carry tcgen05 / TMA code for sm_100a (no legacy mma path)."""
    from wsinsight_b200 import _cabi

    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", str(_cabi.LIB_PATH)], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
    assert "UTCHMMA" in sass or "UTCMMA" in sass, "no tcgen05.mma in SASS"
    assert "UTMALDG" in sass, "no TMA loads in SASS"
    assert "LDTM" in sass, "no tcgen05.ld in SASS"
    assert "HMMA." not in sass.replace("UTCHMMA", ""), "legacy mma.sync present"


def test_no_cuda_means_loud_failure():
    from wsinsight_b200 import _cabi
    from wsinsight_b200.patchlib import segment_tissue

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        segment_tissue(np.zeros((16, 16, 3), np.uint8))
    with pytest.raises(RuntimeError):
        _cabi.require_cuda()


def test_product_never_imports_oracle():
    for p in (ROOT / "wsinsight_b200").rglob("*.py"):
        txt = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), p


def test_lower_resnet34_frozen_and_unfrozen(tmp_path):
    from oracle.forward import make_resnet34
    from wsinsight_b200.modellib.lowering import lower_torchscript

    m = make_resnet34(2, 0)
    for frozen in (True, False):
        ts = torch.jit.script(m)
        if frozen:
            ts = torch.jit.freeze(ts)
        p = tmp_path / f"m{int(frozen)}.pt"
        torch.jit.save(ts, str(p))
        nodes = lower_torchscript(p, 224)
        convs = [n for n in nodes if n.kind == "conv"]
        assert len(convs) == 36
        assert sum(n.relu for n in convs) == 33
        assert sum(n.residual is not None for n in convs) == 16
        assert [n.kind for n in nodes].count("pool") == 1 and nodes[-1].kind == "head"
        assert nodes[-1].fc_w.shape == (2, 512)
        # BN folded: the first conv carries a bias
        assert convs[0].bias is not None and convs[0].shape == (64, 112, 112)
        macs = sum(n.weight.numel() * n.shape[1] * n.shape[2] for n in convs) + nodes[-1].fc_w.numel()
        assert abs(macs / 1e6 - 3663) < 10     # 3.663 GMAC


def test_lowering_rejects_unknown_graphs():
    from wsinsight_b200.errors import UnknownArchitectureError
    from wsinsight_b200.modellib.lowering import lower_torchscript

    class Odd(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.c = torch.nn.Conv2d(3, 8, 3)

        def forward(self, x):
            return torch.tanh(self.c(x)).mean((2, 3))

    with pytest.raises(UnknownArchitectureError):
        lower_torchscript(torch.jit.script(Odd().eval()), 32)

    class RawAndActivated(torch.nn.Module):
        """a conv output read both raw (by a second conv) and through ReLU: fusing the ReLU into the producer
        would hand the activated tensor to the raw reader"""
        def __init__(self):
            super().__init__()
            self.c1 = torch.nn.Conv2d(3, 8, 3, padding=1)
            self.c2 = torch.nn.Conv2d(8, 8, 3, padding=1)
            self.fc = torch.nn.Linear(8, 2)

        def forward(self, x):
            y = self.c1(x)
            z = self.c2(y) + torch.relu(y)
            return self.fc(torch.flatten(torch.nn.functional.adaptive_avg_pool2d(torch.relu(z), (1, 1)), 1))

    with pytest.raises(UnknownArchitectureError, match="un-activated"):
        lower_torchscript(torch.jit.script(RawAndActivated().eval()), 32)

    class PreAddRead(torch.nn.Module):
        """the pre-add value of the only fusable conv feeds another conv as well"""
        def __init__(self):
            super().__init__()
            self.c1 = torch.nn.Conv2d(3, 8, 3, padding=1)
            self.c2 = torch.nn.Conv2d(8, 8, 3, padding=1)
            self.c3 = torch.nn.Conv2d(8, 8, 3, padding=1)
            self.fc = torch.nn.Linear(8, 2)

        def forward(self, x):
            a = torch.relu(self.c1(x))
            y = self.c2(a)
            z = torch.relu(y + a) + torch.relu(self.c3(y))
            return self.fc(torch.flatten(torch.nn.functional.adaptive_avg_pool2d(z, (1, 1)), 1))

    with pytest.raises(UnknownArchitectureError):
        lower_torchscript(torch.jit.script(PreAddRead().eval()), 32)


def test_transform_config_parsing():
    from wsinsight_b200.modellib.models import ModelConfiguration
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from oracle.forward import BREAST_RESNET34_CONFIG

    cfg = ModelConfiguration.from_dict(BREAST_RESNET34_CONFIG)
    assert cfg.class_names == ["Other", "Tumor"] and cfg.patch_size_pixels == 350
    t = make_compose_from_transform_config(cfg.transform)
    assert t.size == 224 and t.mean == [0.7238, 0.5716, 0.6779] and t.out_size(350) == 224
    with pytest.raises(KeyError):
        make_compose_from_transform_config([{"name": "Scale"}, {"name": "ToTensor"}])
    with pytest.raises(ValueError):
        ModelConfiguration.from_dict({"architecture": "x"})


def test_slide_scalars_match_reference_arithmetic():
    from oracle import patchgrid
    from wsinsight_b200.patchlib.pipeline import slide_scalars

    class S:
        dimensions = (100000, 80000)
        mpp = 0.5

    got = slide_scalars(S, 350, 0.25, (2048, 2048), 200 ** 2, 190 ** 2)
    want = patchgrid.slide_params((100000, 80000), (2048, 1638), 0.5, 350, 0.25)
    for k in ("patch_size", "half_patch_size", "min_object_size_px", "min_hole_size_px"):
        assert got[k] == want[k], k


def test_patch_file_roundtrip(tmp_path):
    from wsinsight_b200.modellib.data import _read_patch_coords
    from wsinsight_b200.patchlib.io import load_patch_file, save_hdf5, square_polygons

    coords = np.array([[0, 0], [0, 350], [350, 0]], dtype=np.int64)
    polys = square_polygons(coords, 350)
    assert polys.shape == (15, 2) and polys[1].tolist() == [349.0, 0.0] and polys[4].tolist() == [0.0, 0.0]
    p = save_hdf5(tmp_path / "patches" / "s.h5", coords, polys, np.array([12, 12], np.int32), 350, 0.25,
                  slide_path="/x/s.svs", slide_mpp=0.25, slide_width=4096, slide_height=4096)
    d = load_patch_file(tmp_path / "patches" / "s.h5")
    assert d["slide_path"] == "/x/s.svs" and d["patch_size"] == 350 and d["coords"].dtype == np.int32
    xywh = _read_patch_coords(p)
    assert xywh.tolist() == [[0, 0, 350, 350], [0, 350, 350, 350], [350, 0, 350, 350]]
    with pytest.raises(ValueError):
        save_hdf5(tmp_path / "bad.h5", np.zeros((3, 3)), None, None, 1, 1.0)


def test_csv_writer_matches_reference_columns(tmp_path, golden_purple):
    from wsinsight_b200.modellib.run_inference import write_slide_csv

    g = golden_purple["breast-tumor-resnet34.tcga-brca"]
    c = np.array(g["coords"])
    xywh = np.concatenate([c, np.full_like(c, g["patch_size"])], 1)
    write_slide_csv(tmp_path / "purple.csv", xywh, np.tile([0.25, 0.75], (len(c), 1)).astype(np.float32),
                    ["Other", "Tumor"])
    lines = (tmp_path / "purple.csv").read_text().splitlines()
    assert lines[0].split(",") == g["columns"]
    assert lines[1].startswith("0,0,350,350,0.25,0.75") and len(lines) == 145


def test_shard_ranges_cover_everything():
    from wsinsight_b200.modellib.run_inference import shard_range

    for n in (0, 1, 7, 260947):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            assert max(h - l for l, h in r) - min(h - l for l, h in r) <= 1


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from wsinsight_b200.modellib.run_inference import gather_probs, shard_range
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank, n, K = dist.get_rank(), 7, 2
full = torch.arange(n * K, dtype=torch.float32).reshape(n, K)
lo, hi = shard_range(n, rank, 2)
out = gather_probs(full[lo:hi].clone(), n, rank, 2)
if rank == 0:
    assert torch.equal(out, full), out
else:
    assert out is None
dist.barrier()
print("ok", rank)
"""


def test_gather_world_size_2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_GLOO_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), str(ROOT), port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs


def test_result_path_selection(monkeypatch):
    """Multi-GPU result path: NCCL gather by default, peer stores on request, anything else is rejected."""
    from wsinsight_b200.modellib.run_inference import _result_path

    monkeypatch.delenv("WSI_B200_RESULT_PATH", raising=False)
    assert _result_path() == "nccl" and _result_path("peer") == "peer" and _result_path("NCCL") == "nccl"
    monkeypatch.setenv("WSI_B200_RESULT_PATH", "peer")
    assert _result_path() == "peer" and _result_path("nccl") == "nccl"
    with pytest.raises(ValueError):
        _result_path("smoke-signals")
