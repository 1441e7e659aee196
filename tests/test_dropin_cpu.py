"""CPU: the host-side mirror is a drop-in for the reference classes, and libmde.so exports the C-ABI."""
import os
import re
import sys

import pytest
import torch

import monodepth_estimation_b200 as mde
from helpers import ROOT

REF = "/root/reference"


def test_library_exports_every_declared_symbol():
    from monodepth_estimation_b200 import _lib

    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "mde.h")).read()
    declared = set(re.findall(r"\b(mde_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/mde.h but not exported by libmde.so"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.mde_version() >= 100


@pytest.mark.parametrize("cfg", [(1, 224, 224, False), (2, 518, 518, True), (32, 1024, 1024, False), (32, 1024, 1024, True), (3, 256, 320, True)])
def test_workspace_bytes_matches_the_python_workspace(cfg):
    """mde_workspace_bytes(cfg) is what a C/C++ host sizes its scratch with; engine.Workspace allocates exactly that."""
    from monodepth_estimation_b200 import ops
    from monodepth_estimation_b200.engine import Workspace

    B, H, W, enh = cfg
    ws = Workspace(torch.device("meta"), B, H, W, enh)
    assert ops.workspace_bytes(B, H, W, enh) == ws.nbytes
    with pytest.raises(RuntimeError):
        ops.workspace_bytes(0, 224, 224, False)


def test_no_cpu_fallback_and_train_mode_raises():
    m = mde.DepthAnything(pretrained=False)
    with pytest.raises(RuntimeError):
        m(torch.rand(1, 3, 224, 224))  # train mode
    m.eval()
    with pytest.raises(RuntimeError):
        m(torch.rand(1, 3, 224, 224))  # CPU: the product path has no fallback
    with pytest.raises(RuntimeError):
        mde.DepthAnything(pretrained=True)  # the ImageNet checkpoint is not in the local cache, and nothing is ever downloaded


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "monodepth-estimation_b200")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference sources only exist in the build container")
def test_state_dict_and_init_identical_to_reference():
    sys.path.insert(0, REF)
    try:
        from models.enhanced_model import EnhancedDepthAnything as RE
        from models.model import DepthAnything as RS
        from models.refinement import MultiScaleDepthFusion as RF
        from models.refinement import TextureAwareRefinementModule as RT
    finally:
        sys.path.remove(REF)
    pairs = [(lambda: mde.DepthAnything(pretrained=False), lambda: RS(pretrained=False)),
             (lambda: mde.EnhancedDepthAnything(pretrained=False), lambda: RE(pretrained=False)),
             (mde.TextureAwareRefinementModule, RT), (mde.MultiScaleDepthFusion, RF)]
    for mine, ref in pairs:
        torch.manual_seed(7)
        a = mine().state_dict()
        torch.manual_seed(7)
        b = ref().state_dict()
        assert list(a) == list(b)
        for k in a:
            assert a[k].shape == b[k].shape and torch.equal(a[k], b[k]), k
    m = mde.EnhancedDepthAnything(pretrained=False)
    assert sum(p.numel() for p in m.encoder.parameters()) > 8e7 and len(list(m.decoder.parameters())) > 0  # train/mars_adapter.py:49-50


def test_pretrained_default_loads_the_cached_imagenet_vit(tmp_path, monkeypatch):
    """The reference's default constructor (pretrained=True, encoder.py:21,32) works when torchvision's checkpoint is in the local
    torch.hub cache: the mirror ViT then holds exactly those tensors (the 768->1000 head is dropped, encoder.py:38)."""
    from torchvision.models import ViT_B_16_Weights
    from torchvision.models.vision_transformer import vit_b_16

    torch.manual_seed(3)
    tv = vit_b_16(weights=None)
    ck = tmp_path / "hub" / "checkpoints"
    ck.mkdir(parents=True)
    torch.save(tv.state_dict(), ck / os.path.basename(ViT_B_16_Weights.IMAGENET1K_V1.url))
    monkeypatch.setattr(torch.hub, "get_dir", lambda: str(tmp_path / "hub"))
    m = mde.DepthAnything()  # pretrained=True
    got = m.encoder.encoder.state_dict()
    want = {k: v for k, v in tv.state_dict().items() if not k.startswith("heads.")}
    assert set(got) == set(want)
    assert all(torch.equal(got[k], want[k]) for k in want)


def test_stale_library_is_not_silently_reused(monkeypatch):
    """if the build fails while the sources are newer than libmde.so, load() must raise instead of binding the old binary."""
    from monodepth_estimation_b200 import _lib
    from monodepth_estimation_b200 import build as B

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(B, "build", lambda *a, **k: (_ for _ in ()).throw(RuntimeError("nvcc failed")))
    monkeypatch.setattr(B, "is_stale", lambda: True)
    with pytest.raises(_lib.MdeError, match="no up-to-date binary"):
        _lib.load()
    monkeypatch.setattr(B, "is_stale", lambda: False)  # binary matches the sources (e.g. a box without nvcc): fine
    assert _lib.load().mde_version() >= 100


def test_no_load_of_predecessor_data_is_scheduled_above_the_pdl_wait():
    """Every kernel runs under programmatic dependent launch: whatever it reads of its predecessor's output must come after
    griddepcontrol.wait (SASS: ACQBULK).  The compiler is free to move `__ldg` / const __restrict__ loads above that instruction —
    it did, in the LayerNorm kernel, once code was placed before the wait (wrong activations under PDL, correct in isolation) — so
    the built library is checked: before ACQBULK a kernel may only touch shared memory (generic LD.E / ATOMS / SYNCS), issue the
    volatile parameter loads of ptx.cuh's ld_param_now_f4 (LDG...STRONG.SYS), or be a kernel known to fetch its weights there."""
    import shutil
    import subprocess

    from monodepth_estimation_b200 import _lib as L

    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    L.load()
    sass = subprocess.run([cuobjdump, "-sass", L.SO_PATH], capture_output=True, text=True, check=True).stdout
    # kernels that fetch PARAMETERS before the wait, by design, and the only opcodes they may use for it: the opt-in fused decoder
    # tail stages its three layers' weights with read-only loads; conv_halo requests its 3x3 weight set by TMA
    weights_before_wait = {"decoder_tail_kernel": ("LDG",), "conv_halo_kernel": ("UTMALDG",)}
    offenders, kernels, fn, waited = [], 0, None, False
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn, waited = m.group(1), False
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", line)
        if not m or fn is None:
            continue
        ins = m.group(2)
        if "ACQBULK" in ins:
            waited, kernels = True, kernels + 1
        elif not waited and re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0] in ("LDG", "STG", "RED", "ATOM", "ATOMG", "UTMALDG", "UBLKCP"):
            op = re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0]
            if "STRONG.SYS" in ins or any(k in fn and op in ops for k, ops in weights_before_wait.items()):
                continue
            offenders.append((fn, m.group(1), ins.strip()))
    # kernels without ACQBULK never reach `waited`; their instructions are not "before the wait", so drop them
    with_wait = set()
    fn = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
        elif "ACQBULK" in line and fn:
            with_wait.add(fn)
    offenders = [o for o in offenders if o[0] in with_wait]
    assert kernels >= 60, kernels  # the library's kernels do carry the wait (guards against a silent change of the mnemonic)
    assert not offenders, offenders[:8]
