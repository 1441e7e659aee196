"""GPU: the backward kernels of the ViT encoder block (csrc/backward.cu + training.py) against torch autograd in
fp32 on the same seeded inputs.  Operands are bf16 with fp32 accumulation, so gradients carry bf16-level error: every test
states its tolerance."""
import pytest
import torch
import torch.nn.functional as F

from helpers import max_abs, rel_l2

pytestmark = pytest.mark.gpu

BF16 = torch.bfloat16
# the torch side is the fp32 reference: no TF32 in its convolutions / matmuls
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def _mods():
    from monodepth_estimation_b200 import _lib as L
    from monodepth_estimation_b200 import ops

    return L, ops


def _rand(*shape, seed=0, scale=1.0, dtype=torch.float32):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).cuda()


@pytest.mark.parametrize("R,C,f32", [(1576, 768, True), (197, 2304, False), (70, 40, True), (1576, 3072, False)])
def test_transpose_cast_pad(R, C, f32):
    L, ops = _mods()
    x = _rand(R, C, seed=1, dtype=torch.float32 if f32 else BF16)
    ldt = (R + 63) // 64 * 64
    t, rm = ops.transpose_bf16(x, ldt=ldt, want_rm=True)
    assert torch.equal(rm, x.to(BF16))
    assert torch.equal(t[:, :R], x.to(BF16).t())
    assert float(t[:, R:].float().abs().sum()) == 0.0  # the K padding of a dW GEMM must be exact zeros


def test_colsum_matches_sum_and_accumulates():
    L, ops = _mods()
    x = _rand(1576, 2304, seed=2, dtype=BF16)
    ref = x.double().sum(0)
    out = ops.colsum_bf16(x)
    assert max_abs(out, ref) < 2e-3
    again = ops.colsum_bf16(x)
    assert torch.equal(out, again)  # fixed summation order
    ops.colsum_bf16(x, out=out, accumulate=True)
    assert max_abs(out, 2 * ref) < 4e-3


def test_gelu_forward_and_backward():
    L, ops = _mods()
    u = _rand(1576, 3072, seed=3, scale=2.0, dtype=BF16)
    dy = _rand(1576, 3072, seed=4, dtype=BF16)
    uf = u.float().requires_grad_(True)
    y = F.gelu(uf)
    y.backward(dy.float())
    assert rel_l2(ops.gelu_bf16(u).float(), y.detach()) < 3e-3
    assert rel_l2(ops.gelu_bwd_bf16(u, dy).float(), uf.grad) < 3e-3


@pytest.mark.parametrize("rows,resid", [(1576, True), (197, False), (65, True)])
def test_layernorm_backward(rows, resid):
    L, ops = _mods()
    x = (_rand(rows, 768, seed=5) * (0.5 + torch.rand(rows, 1, device="cuda")) + 0.2 * _rand(rows, 1, seed=6)).contiguous()
    gamma, beta = 1.0 + 0.2 * _rand(768, seed=7), 0.1 * _rand(768, seed=8)
    dy = _rand(rows, 768, seed=9, dtype=BF16)
    dres = _rand(rows, 768, seed=10) if resid else None
    xr, gr, br = x.clone().requires_grad_(True), gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    F.layer_norm(xr, (768,), gr, br, 1e-6).backward(dy.float())
    dx, dg, db = ops.layernorm_bwd(x, dy, gamma, 1e-6, dres=dres)
    want = xr.grad + (dres if resid else 0.0)
    assert rel_l2(dx, want) < 1e-5, rel_l2(dx, want)
    assert rel_l2(dg, gr.grad) < 1e-5 and rel_l2(db, br.grad) < 1e-5
    dg2 = dg.clone()
    ops.layernorm_bwd(x, dy, gamma, 1e-6, dgamma=dg2, dbeta=db.clone(), accumulate=True)
    assert rel_l2(dg2, 2 * gr.grad) < 1e-5


@pytest.mark.parametrize("B", [1, 3])
def test_attention_backward(B):
    """dq | dk | dv of softmax(q k^T / 8) v per (image, head), against autograd on the bf16-rounded operands."""
    L, ops = _mods()
    M = B * 197
    h1 = _rand(M, 768, seed=11, dtype=BF16)
    w, b = _rand(2304, 768, seed=12, scale=768 ** -0.5, dtype=BF16), _rand(2304, seed=13, scale=0.1)
    q = torch.empty(M, 768, device="cuda", dtype=BF16)
    k = torch.empty_like(q)
    vt = torch.zeros(B * 12 * 64 * L.VT_PITCH, device="cuda", dtype=BF16)
    ops.gemm(h1, w, b, L.EPI_QKV_SPLIT, q, k, vt)  # the layouts exactly as the forward writes them
    dout = _rand(M, 768, seed=14, dtype=BF16)
    dqkv = ops.vit_attention_bwd(q, k, vt, dout, B)
    # reference: unscaled q (= 8 * stored q), k, v as leaves
    qf = (q.float() * 8.0).view(B, 197, 12, 64).permute(0, 2, 1, 3).clone().requires_grad_(True)
    kf = k.float().view(B, 197, 12, 64).permute(0, 2, 1, 3).clone().requires_grad_(True)
    vf = vt.view(B, 12, 64, L.VT_PITCH)[..., :197].float().transpose(-1, -2).clone().requires_grad_(True)
    o = torch.softmax(qf @ kf.transpose(-1, -2) * 0.125, -1) @ vf
    o.backward(dout.float().view(B, 197, 12, 64).permute(0, 2, 1, 3))
    back = lambda t: t.permute(0, 2, 1, 3).reshape(M, 768)
    got = dqkv.float()
    assert rel_l2(got[:, :768], back(qf.grad)) < 6e-3, rel_l2(got[:, :768], back(qf.grad))
    assert rel_l2(got[:, 768:1536], back(kf.grad)) < 6e-3
    assert rel_l2(got[:, 1536:], back(vf.grad)) < 6e-3
    assert torch.equal(dqkv, ops.vit_attention_bwd(q, k, vt, dout, B))  # deterministic


@pytest.mark.parametrize("B", [2, 8])
def test_vit_block_backward_matches_autograd(B):
    """One torchvision EncoderBlock (tv:84-120): forward through the C-ABI kernels, backward through training.ViTBlockTrainer,
    against torch autograd of the same block in fp32.  B = 8 is config 5's per-GPU batch (M = 1576: the CTA-pair GEMM)."""
    from torchvision.models.vision_transformer import EncoderBlock

    from monodepth_estimation_b200.training import ViTBlockTrainer

    torch.manual_seed(21)
    blk = EncoderBlock(12, 768, 3072, 0.0, 0.0).cuda().float()
    with torch.no_grad():
        for n, p_ in blk.named_parameters():
            if n.startswith("ln_"):
                p_.add_(0.1 * torch.randn_like(p_))
            elif n.endswith("bias"):
                p_.normal_(0, 0.05)
    M = B * 197
    x = _rand(M, 768, seed=22)
    dout = _rand(M, 768, seed=23, scale=0.1)
    xr = x.view(B, 197, 768).clone().requires_grad_(True)
    y = blk(xr)
    y.backward(dout.view(B, 197, 768))
    tr = ViTBlockTrainer(dict(blk.named_parameters()))
    out = tr.forward(x)
    assert rel_l2(out, y.detach().reshape(M, 768)) < 4e-3
    dx, grads = tr.backward(dout)
    assert rel_l2(dx, xr.grad.reshape(M, 768)) < 1e-2, rel_l2(dx, xr.grad.reshape(M, 768))
    for n, p_ in blk.named_parameters():
        e = rel_l2(grads[n], p_.grad)
        assert grads[n].shape == p_.grad.shape and e < 2e-2, (n, e)


def test_vit_encoder_backward_matches_autograd():
    """The whole torchvision ViT-B/16 as encoder.py uses it (hooks on blocks 2/5/8/11, final ln discarded): gradients of every
    parameter that autograd reaches, for a loss that reads taps 8 and 11 only — what ('disp', 0) depends on (decoder.py:97-166)."""
    import torchvision

    from monodepth_estimation_b200.training import HOOKS, ViTEncoderTrainer

    torch.manual_seed(31)
    B = 2
    vit = torchvision.models.vit_b_16(weights=None).cuda().float()
    vit.heads = torch.nn.Identity()
    with torch.no_grad():
        vit.class_token.normal_(0, 0.02)
        for n, p_ in vit.named_parameters():
            if ".ln" in n or "ln_" in n:
                p_.add_(0.05 * torch.randn_like(p_))
    feats = []
    hooks = [vit.encoder.layers[i].register_forward_hook(lambda m, a, o: feats.append(o)) for i in HOOKS]
    x224 = _rand(B, 3, 224, 224, seed=32)
    vit(x224)
    for h in hooks:
        h.remove()
    dt = [None, None, _rand(B * 197, 768, seed=33, scale=0.1), _rand(B * 197, 768, seed=34, scale=0.1)]
    loss = (feats[2].reshape(-1, 768) * dt[2]).sum() + (feats[3].reshape(-1, 768) * dt[3]).sum()
    loss.backward()

    patches = x224.view(B, 3, 14, 16, 14, 16).permute(0, 2, 4, 1, 3, 5).reshape(B * 196, 768).to(BF16).contiguous()
    params = {n: p_ for n, p_ in vit.named_parameters()}
    tr = ViTEncoderTrainer(params)
    taps = tr.forward(patches)
    for i in range(4):
        assert rel_l2(taps[i], feats[i].detach().reshape(-1, 768)) < 6e-3, i
    grads = tr.backward(dt)
    worst = ("", 0.0)
    for n, p_ in vit.named_parameters():
        if p_.grad is None:
            assert n not in grads, n
            continue
        e = rel_l2(grads[n], p_.grad)
        worst = max(worst, (n, e), key=lambda t: t[1])
        assert grads[n].shape == p_.grad.shape and e < 3e-2, (n, e)
    print("worst parameter gradient", worst)


# --------------------------------------------------------------------------- train-mode decoder path, loss, Adam (csrc/train_ops.cu)
@pytest.mark.parametrize("pad", ["reflect", "zero"])
def test_im2col_gemm_is_the_conv_and_col2im_is_its_adjoint(pad):
    L, ops = _mods()
    B, h, w, C, Co = 2, 14, 15, 64, 128
    x = _rand(B, h, w, C, seed=41)
    wt = _rand(Co, C, 3, 3, seed=42, scale=0.05)
    mode = L.PAD_REFLECT if pad == "reflect" else L.PAD_ZERO
    col, colT = ops.im2col3x3(x, mode)
    n = B * h * w
    assert torch.equal(colT[:, :n], col.t()) and float(colT[:, n:].float().abs().sum()) == 0.0
    xn = x.permute(0, 3, 1, 2).clone().requires_grad_(True)
    xp = F.pad(xn, (1, 1, 1, 1), mode="reflect") if pad == "reflect" else F.pad(xn, (1, 1, 1, 1))
    ref = F.conv2d(xp, wt)
    wm = wt.permute(0, 2, 3, 1).reshape(Co, 9 * C)  # k = (ky*3+kx)*C + c
    got = (col.float() @ wm.t()).view(B, h, w, Co).permute(0, 3, 1, 2)
    assert rel_l2(got, ref.detach()) < 4e-3
    dy = _rand(B, Co, h, w, seed=43)
    ref.backward(dy)
    dxcol = (dy.permute(0, 2, 3, 1).reshape(n, Co) @ wm).contiguous()  # fp32 here: isolates the col2im fold
    dx = ops.col2im3x3(dxcol, B, h, w, C, mode)
    assert rel_l2(dx.permute(0, 3, 1, 2), xn.grad) < 1e-5


def test_bn_elu_train_forward_backward_and_running_stats():
    L, ops = _mods()
    N, C = 1680, 128
    z = (_rand(N, C, seed=44) * (0.5 + torch.rand(1, C, device="cuda")) + _rand(1, C, seed=45)).contiguous()
    bn = torch.nn.BatchNorm2d(C).cuda().train()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.normal_(0, 0.2)
        bn.running_mean.normal_(0, 0.1)
        bn.running_var.uniform_(0.5, 1.5)
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    zr = z.t().reshape(1, C, N, 1).clone().requires_grad_(True)
    y = F.elu(bn(zr))
    dy = _rand(N, C, seed=46)
    y.backward(dy.t().reshape(1, C, N, 1))
    got, mean, rstd = ops.bn_elu_train_fwd(z, C, bn.weight.detach(), bn.bias.detach(), run_mean=rm, run_var=rv)
    assert rel_l2(got, y.detach().view(C, N).t()) < 1e-5
    assert rel_l2(rm, bn.running_mean) < 1e-6 and rel_l2(rv, bn.running_var) < 1e-5
    dz, dg, db = ops.bn_elu_train_bwd(z, C, mean, rstd, bn.weight.detach(), bn.bias.detach(), dy)
    assert rel_l2(dz, zr.grad.view(C, N).t()) < 1e-4
    assert rel_l2(dg, bn.weight.grad) < 1e-4 and rel_l2(db, bn.bias.grad) < 1e-4


@pytest.mark.parametrize("hs,ws,hd,wd,align,C", [(14, 15, 28, 30, True, 128), (28, 30, 528, 528, False, 1), (28, 30, 130, 97, False, 1), (14, 15, 28, 30, False, 8)])
def test_bilinear_forward_and_adjoint(hs, ws, hd, wd, align, C):
    L, ops = _mods()
    B = 2
    x = _rand(B, hs, ws, C, seed=47)
    xr = x.permute(0, 3, 1, 2).clone().requires_grad_(True)
    ref = F.interpolate(xr, size=(hd, wd), mode="bilinear", align_corners=align)
    got = ops.bilinear_nhwc(x, hd, wd, align)
    assert rel_l2(got.permute(0, 3, 1, 2), ref.detach()) < 1e-5
    dd = _rand(B, hd, wd, C, seed=48)
    ref.backward(dd.permute(0, 3, 1, 2))
    ds = ops.bilinear_nhwc_bwd(dd, hs, ws, align)
    assert rel_l2(ds.permute(0, 3, 1, 2), xr.grad) < 1e-5


def test_instnorm_backward_and_grid_adjoint():
    L, ops = _mods()
    B, G, C = 3, 210, 256
    x = _rand(B, G, C, seed=49) * 2 + 0.3
    xr = x.permute(0, 2, 1).reshape(B, C, 14, 15).clone().requires_grad_(True)
    y = F.instance_norm(xr, eps=1e-5)
    dy = _rand(B, G, C, seed=50)
    y.backward(dy.permute(0, 2, 1).reshape(B, C, 14, 15))
    assert rel_l2(ops.instnorm_f32(x).permute(0, 2, 1).reshape(B, C, 14, 15), y.detach()) < 1e-5
    assert rel_l2(ops.instnorm_bwd_f32(x, dy).permute(0, 2, 1).reshape(B, C, 14, 15), xr.grad) < 1e-4
    dgrid = _rand(B, G, 768, seed=51)
    dtok = ops.tokens_to_grid_bwd(dgrid, B, 0, G).view(B, 197, 768)
    assert torch.equal(dtok, dgrid[:, :197])  # 14x15 grid: CLS kept at cell 0, cells 197..209 are zero padding (encoder.py:90-102)
    dgrid2 = _rand(B, 196, 768, seed=52)
    dtok2 = ops.tokens_to_grid_bwd(dgrid2, B, 1, 196).view(B, 197, 768)
    assert float(dtok2[:, 0].abs().sum()) == 0.0 and torch.equal(dtok2[:, 1:], dgrid2)


def test_depth_loss_matches_compute_losses():
    """MarsDomainAdapter.compute_losses (train/mars_adapter.py:209-247) restated in torch with the rand draw passed in."""
    L, ops = _mods()
    B, H, W = 2, 96, 128
    disp = torch.sigmoid(_rand(B, 1, H, W, seed=53)).contiguous()
    img = torch.rand(B, 3, H, W, device="cuda")
    rnd = torch.rand(B, 1, H, W, device="cuda")
    d = disp.clone().requires_grad_(True)
    gx = (d[:, :, :, :-1] - d[:, :, :, 1:]).abs() * torch.exp(-(img[:, :, :, :-1] - img[:, :, :, 1:]).abs().mean(1, keepdim=True))
    gy = (d[:, :, :-1, :] - d[:, :, 1:, :]).abs() * torch.exp(-(img[:, :, :-1, :] - img[:, :, 1:, :]).abs().mean(1, keepdim=True))
    smooth = 0.001 * (gx.mean() + gy.mean())
    depth = 1.0 * (d - d.detach() * (0.9 + 0.2 * rnd)).abs().mean()
    (depth + smooth).backward()
    losses, dd = ops.depth_loss(disp, img, rnd, 1.0, 0.001)
    assert abs(float(losses[1]) - float(depth)) < 1e-6 * float(depth) + 1e-9
    assert abs(float(losses[2]) - float(smooth)) < 1e-5 * float(smooth) + 1e-10
    assert abs(float(losses[0]) - float(depth + smooth)) < 1e-6 * float(depth + smooth)
    assert rel_l2(dd, d.grad) < 1e-5


def test_adam_step_matches_torch():
    L, ops = _mods()
    p0, lr = _rand(100003, seed=54), 1e-4
    p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([p], lr=lr)
    mine, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    for step in range(1, 4):
        g = _rand(100003, seed=60 + step)
        p.grad = g.clone()
        opt.step()
        ops.adam_step(mine, g, m, v, lr, step)
    assert max_abs(mine, p.detach()) < 5e-7  # one ulp at |p| ~ 3


@pytest.mark.parametrize("scale255", [False, True])
def test_input_norm_bwd_matches_autograd(scale255):
    """encoder.py:129-137: gradient of the six BatchNorm2d(3) scalars in front of the patch embedding, through the range heuristic
    (both branches: inputs already in [0, 1] -> 2x - 1; inputs up to 255 -> / 255 first)."""
    L, ops = _mods()
    B = 3
    g = torch.Generator().manual_seed(71)
    raw = torch.rand(B, 3, 224, 224, generator=g).cuda() * (255.0 if scale255 else 1.0)
    mean, var = raw.mean((0, 2, 3)), raw.var((0, 2, 3), unbiased=False)
    # weight / bias that keep the normalised values inside the branch under test
    w = torch.tensor([0.11, 0.09, 0.1]).cuda() * (255.0 if scale255 else 1.0)
    b = torch.tensor([0.5, 0.45, 0.55]).cuda() * (255.0 if scale255 else 1.0)
    bn = torch.cat([w, b, mean, var]).contiguous()
    scratch = torch.empty(L.RANGE_SCRATCH_FLOATS, device="cuda")
    affine = torch.empty(6, device="cuda")
    raw224 = torch.empty(B, 3, 224, 224, device="cuda")
    patches = torch.empty(B * 196, 768, device="cuda", dtype=torch.bfloat16)
    ops.input_stage(raw, bn, 1e-5, scratch, affine, raw224, patches)
    dP = _rand(B * 196, 768, seed=72)
    got = ops.input_norm_bwd(dP, raw224, affine, bn, 1e-5)
    wt, bt = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = torch.nn.functional.batch_norm(raw, None, None, wt, bt, True, 0.1, 1e-5)
    assert (float(y.max()) > 1.0) == scale255
    if scale255:
        y = y / 255.0
    assert float(y.min()) >= 0.0 and float(y.max()) <= 1.0
    y = y * 2.0 - 1.0
    rows = y.view(B, 3, 14, 16, 14, 16).permute(0, 2, 4, 1, 3, 5).reshape(B * 196, 768)  # (image, patch) x (c, ky, kx)
    assert rel_l2(patches.float(), rows.detach()) < 5e-3
    (rows * dP).sum().backward()
    want = torch.cat([wt.grad, bt.grad])
    assert rel_l2(got, want) < 1e-4, (got, want)


@pytest.mark.parametrize("B,H,W", [(2, 176, 240), (8, 518, 518)], ids=["b2_176x240", "config5_b8_518"])
def test_training_step_matches_train_oracle(B, H, W):
    """DepthTrainer (train/mars_adapter.py:117-153 on the device) against oracle/train_oracle.py — itself pinned to the real
    reference classes in train() mode (tests/test_train_oracle.py): losses, every parameter gradient, BatchNorm running
    statistics, and the parameters after one Adam step.  The second case is BASELINE.json configs[4] (B=8 at 518x518)."""
    import monodepth_estimation_b200 as mde
    from helpers import ledger_add
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from monodepth_estimation_b200.training import DepthTrainer
    from oracle import train_oracle as T

    case = f"train:b{B}_{H}x{W}"
    m = mde.DepthAnything(pretrained=False)
    sd0 = synth_state_dict(m.state_dict(), 0)
    # F.instance_norm (encoder.py:167) divides by the spread of a channel ACROSS grid positions.  With the stock random init every
    # token is nearly the same vector (pos_embedding std 0.02), that spread is ~3 % of the values, and the bf16 rounding of the ViT
    # (4e-3) comes out of the normalisation as a 6 % error of the features whatever the kernels do.  Position embeddings of unit
    # scale (as trained ones are) make the comparison measure the implementation instead: same weights on both sides.
    sd0["encoder.encoder.encoder.pos_embedding"] = sd0["encoder.encoder.encoder.pos_embedding"] * 50.0
    m.load_state_dict(sd0)
    x = synth_images(B, H, W, 5)
    md = {"source": "rover", "height": 1.6}
    rnd = torch.rand(B, 1, H, W, generator=torch.Generator().manual_seed(9))
    sd_gpu = {k: v.cuda() for k, v in sd0.items()}
    want_losses, want_grads, want_buffers = T.loss_and_grads(sd_gpu, x.cuda(), md, rnd.cuda())

    m = m.cuda()
    tr = DepthTrainer(m, lr=1e-4)
    losses, grads = tr.forward_backward(x.cuda(), md, rnd.cuda())
    got = [float(v) for v in losses.cpu()]
    ledger_add(case, "loss/depth", "rel", abs(got[1] - want_losses["depth"]) / want_losses["depth"], 2e-2, "training step")
    ledger_add(case, "loss/smoothness", "rel", abs(got[2] - want_losses["smoothness"]) / want_losses["smoothness"], 6e-2, "training step")
    assert abs(got[1] - want_losses["depth"]) < 2e-2 * want_losses["depth"], (got, want_losses)
    # the smoothness term sums |differences of neighbouring disparities|: bf16-level noise (3e-3) on a map whose neighbours differ by
    # a few per cent is amplified accordingly
    assert abs(got[2] - want_losses["smoothness"]) < 6e-2 * want_losses["smoothness"], (got, want_losses)
    skipped = set()
    errs = []
    for k, gw in want_grads.items():
        if k in skipped:
            continue
        assert k in grads, k
        a, b = grads[k].reshape(gw.shape), gw
        if float(b.abs().max()) < 1e-9:  # analytically zero (a bias in front of a normalisation): rounding noise on both sides
            assert float(a.abs().max()) < 1e-4 * max(1.0, float(want_losses["total"])), k
            continue
        e = rel_l2(a, b)
        errs.append((e, k))
    errs.sort(reverse=True)
    print("non-ViT gradient errors", [(k.replace("decoder.convs.", "d."), round(e, 4)) for e, k in errs if "encoder.encoder." not in k])
    print("losses", got, "largest gradient errors", [(k, round(e, 4)) for e, k in errs[:12]], "median", errs[len(errs) // 2][0])
    ledger_add(case, "grad/" + errs[0][1], "rel_l2", errs[0][0], 5e-2, "training step: worst parameter gradient")
    ledger_add(case, "grad/median", "rel_l2", errs[len(errs) // 2][0], 5e-2, "training step: median parameter gradient")
    assert errs[0][0] < 5e-2, errs[:5]
    assert set(grads) <= set(want_grads), set(grads) - set(want_grads)
    # every BatchNorm buffer of the model after ONE train-mode forward, as model.train(); model(x) leaves them in the reference:
    # encoder.normalize + all 13 decoder ConvBlocks (levels 1-3 run forward-only: decoder.py:106-168), num_batches_tracked included
    sd_now = m.state_dict()
    n_stats = 0
    worst = ("", 0.0)
    for k, wb in want_buffers.items():
        if k.endswith(("running_mean", "running_var")):
            e = rel_l2(sd_now[k], wb)
            worst = max(worst, (k, e), key=lambda t: t[1])
            assert e < 2e-2, (k, e)
            n_stats += 1
        elif k.endswith("num_batches_tracked"):
            assert int(sd_now[k]) == int(wb) == 1, k
    assert n_stats == 28, n_stats
    ledger_add(case, "bn/" + worst[0], "rel_l2", worst[1], 2e-2, "training step: worst BatchNorm running statistic")
    print("worst running statistic", worst)
    # ---- one optimizer step (train/mars_adapter.py:47-55,139): Adam, encoder.* at lr/10, decoder.* at lr
    before = {k: v.detach().clone() for k, v in m.state_dict().items()}
    tr.step(x.cuda(), md, rnd.cuda())
    after = m.state_dict()
    ref_params = {k: sd_gpu[k].clone() for k in want_grads}
    T.adam_reference(ref_params, {k: v for k, v in want_grads.items() if k not in skipped}, {}, lambda k: 1e-5 if k.startswith("encoder.") else 1e-4, 1)
    for k in ("decoder.convs.upconv_0_1.conv.0.conv.1.weight", "encoder.encoder.encoder.layers.encoder_layer_11.mlp.3.weight", "encoder.channel_reduce.3.weight",
              "encoder.encoder.conv_proj.weight"):
        lr_k = 1e-5 if k.startswith("encoder.") else 1e-4
        mine, want = after[k] - before[k], ref_params[k] - sd_gpu[k]
        assert float(mine.abs().max()) <= 1.001 * lr_k  # the first Adam step moves every element by at most lr
        # g / (|g| + eps) is a sign: elements whose gradient is within the bf16 error of zero may flip, the rest must agree
        bad = float(((mine - want).abs() > 0.5 * lr_k).float().mean())
        assert bad < 0.03, (k, bad)
    untouched = "decoder.convs.upconv_3_1.conv.0.conv.1.weight"  # no gradient reaches levels 1-3: Adam leaves them alone
    assert torch.equal(after[untouched], before[untouched])


@pytest.mark.parametrize("with_hook", [False, True])
def test_graph_captured_steps_equal_eager_steps(with_hook):
    """DepthTrainer(graph=True) replays CUDA graphs from the third step on; every kernel is deterministic, so four steps (eager,
    capture, replay, replay) must leave exactly the parameters, Adam moments and running statistics of four eager steps, with
    different images each step.  With a gradient hook (here: halve every gradient, standing in for the all-reduce) the backward
    is captured as one graph per stage and the hook runs on a side stream next to the following stage."""
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from monodepth_estimation_b200.training import DepthTrainer

    B, H, W = 2, 176, 240
    md = {"source": "rover"}
    results, calls = [], []

    def halve(part):
        calls.append(len(part))
        for g in part.values():
            g.mul_(0.5)

    hook = halve if with_hook else None
    for graph in (False, True):
        m = mde.DepthAnything(pretrained=False)
        m.load_state_dict(synth_state_dict(m.state_dict(), 0))
        m = m.cuda()
        tr = DepthTrainer(m, lr=1e-3, graph=graph)
        losses = []
        for i in range(4):
            x = synth_images(B, H, W, 30 + i).cuda()
            rnd = torch.rand(B, 1, H, W, generator=torch.Generator().manual_seed(40 + i)).cuda()
            losses.append(tr.step(x, md, rnd, hook).clone())
        torch.cuda.synchronize()
        results.append(({k: v.detach().clone() for k, v in m.state_dict().items()}, torch.stack(losses), tr))
    (sd_e, l_e, tr_e), (sd_g, l_g, tr_g) = results
    assert torch.equal(l_e, l_g), (l_e, l_g)
    assert float((l_e[0] - l_e[3]).abs().max()) > 0  # the steps did something
    for k in sd_e:
        assert torch.equal(sd_e[k], sd_g[k]), k
    assert int(tr_e.step_dev) == int(tr_g.step_dev) == 4
    if with_hook:  # eager: one call per step with everything; graph: one eager step, then len(splits) + 1 parts per step
        assert calls[:4] == [calls[0]] * 4 and len(calls) == 4 + 1 + 3 * (len(tr_g.splits) + 1) and sum(calls[5:9]) == calls[0]
    for k in tr_e.state:
        assert torch.equal(tr_e.state[k][1], tr_g.state[k][1]), k


def test_adam_multi_matches_torch_and_reads_lr_from_the_device():
    """mde_adam_multi: ONE launch over tensors of ragged sizes (vector path, unaligned tails, a 1-element tensor), two
    learning-rate groups read from device memory — so a schedule change between steps takes effect without re-building anything
    (train/mars_adapter.py:47-61)."""
    L, ops = _mods()
    sizes = [100003, 4096, 7, 1, 768 * 3072, 12289]
    ps = [torch.nn.Parameter(_rand(n, seed=100 + i)) for i, n in enumerate(sizes)]
    groups = [i % 2 for i in range(len(sizes))]
    lrs = [1e-5, 1e-4]
    opt = torch.optim.Adam([{"params": [p for p, g in zip(ps, groups) if g == 0], "lr": lrs[0]},
                            {"params": [p for p, g in zip(ps, groups) if g == 1], "lr": lrs[1]}])
    mine = [p.detach().clone() for p in ps]
    gbuf = [torch.empty_like(p) for p in mine]
    ms, vs = [torch.zeros_like(p) for p in mine], [torch.zeros_like(p) for p in mine]
    plan = ops.AdamPlan([(p, g, m, v, grp) for p, g, m, v, grp in zip(mine, gbuf, ms, vs, groups)])
    lr_dev = torch.tensor(lrs, device="cuda")
    step_dev = torch.zeros(1, device="cuda", dtype=torch.int32)
    for step in range(1, 5):
        if step == 3:  # StepLR fires: both groups halve
            for gparam in opt.param_groups:
                gparam["lr"] *= 0.5
            lr_dev.mul_(0.5)
        for i, p in enumerate(ps):
            g = _rand(sizes[i], seed=200 + 10 * step + i)
            p.grad = g.clone()
            gbuf[i].copy_(g)
        opt.step()
        step_dev += 1
        plan.step(lr_dev, step_dev)
    for a, p in zip(mine, ps):
        assert max_abs(a, p.detach()) < 1e-6


def test_inference_after_training_uses_the_new_weights():
    """DepthTrainer rewrites parameters and running statistics through raw pointers, so tensor._version does not
    move; the engine must still re-pack.  infer -> train -> infer must equal a fresh model loaded from the new state_dict."""
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from monodepth_estimation_b200.training import DepthTrainer

    m = mde.DepthAnything(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.cuda().eval()
    x = synth_images(2, 176, 240, 5).cuda()
    md = {"source": "rover"}
    before = m(x, md)[("disp", 0)].clone()
    m(x, md)  # graph replay path is live
    tr = DepthTrainer(m, lr=1e-2)
    rnd = torch.rand(2, 1, 176, 240, generator=torch.Generator().manual_seed(9)).cuda()
    for _ in range(3):  # eager, capture, replay
        m.train()
        tr.step(x, md, rnd)
        m.eval()
        after = m(x, md)[("disp", 0)].clone()
        fresh = mde.DepthAnything(pretrained=False)
        fresh.load_state_dict({k: v.detach().cpu().clone() for k, v in m.state_dict().items()})
        want = fresh.cuda().eval()(x, md)[("disp", 0)]
        assert torch.equal(after, want)
    assert not torch.equal(after, before)
    nbt = m.state_dict()["decoder.convs.init_conv.conv.1.num_batches_tracked"]
    assert int(nbt) == 3 and int(m.state_dict()["encoder.normalize.num_batches_tracked"]) == 3


def test_learning_rate_change_reaches_the_captured_step():
    """the learning rate is read from device memory by the captured Adam launch; lr = 0 must freeze the weights."""
    import monodepth_estimation_b200 as mde
    from monodepth_estimation_b200.synth import synth_images, synth_state_dict
    from monodepth_estimation_b200.training import DepthTrainer

    m = mde.DepthAnything(pretrained=False)
    m.load_state_dict(synth_state_dict(m.state_dict(), 0))
    m = m.cuda()
    tr = DepthTrainer(m, lr=1e-3)
    x = synth_images(2, 176, 240, 6).cuda()
    rnd = torch.rand(2, 1, 176, 240, generator=torch.Generator().manual_seed(3)).cuda()
    key = "decoder.convs.upconv_0_1.conv.0.conv.1.weight"
    for _ in range(3):
        tr.step(x, None, rnd)  # third step replays the graphs
    w3 = m.state_dict()[key].clone()
    tr.set_lr(0.0)
    tr.step(x, None, rnd)
    assert torch.equal(m.state_dict()[key], w3)
    tr.lr_scheduler_step(epoch=5, step_size=5, gamma=0.5, base_lr=1e-3)  # StepLR: 1e-3 * 0.5
    assert abs(tr.lr - 5e-4) < 1e-12
    tr.step(x, None, rnd)
    d = (m.state_dict()[key] - w3).abs().max()
    assert 0 < float(d) <= 5e-4 * 1.5  # moved, by at most ~lr (Adam's per-step bound once the moments are warm: <= lr * (1 + small))
