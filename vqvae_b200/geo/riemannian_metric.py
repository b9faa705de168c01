"""Riemannian (decoder pull-back) edge lengths -- drop-in for the reference's
`src/geo/riemannian_metric.py` (citations are lines of that file).

    L(i,j) = 0.5 * ( |J(z_i) dz| + |J(z_j) dz| ),   J = Jacobian of sigmoid(decoder(z)),  dz = z_j - z_i

The JVPs stay in PyTorch (north star): conv/BN/ReLU stacks are propagated in explicit
forward mode (primal and tangent share one convolution call) with BatchNorm statistics
taken per `batch_size`-edge batch exactly as the reference's loop does (:50-58) -- the
reference CLI leaves the decoder in train mode (build_codebook.py:27-29), so results
depend on batch membership.  Many batches are processed per call by grouping the BN
statistics.  The per-edge reduction + scatter into both CSR directions is the CUDA
kernel `geocb_edge_len_scatter`.  Unknown module types fall back to
`torch.autograd.functional.jvp` per batch (still on the GPU).
"""
from __future__ import annotations

import contextlib
import os
import copy

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import _lib
from .. import device_graph as dg


@torch.no_grad()
def decoder_logits_to_img(logits: torch.Tensor) -> torch.Tensor:
    """Convert decoder logits to image space [0,1] using sigmoid (:7-10)."""
    return torch.sigmoid(logits)


def _is_flat_decoder(decoder) -> bool:
    return hasattr(next(decoder.children()), "in_features")  # :18-20


def _compute_jacobian_vector_product(decoder, z: torch.Tensor, direction: torch.Tensor) -> torch.Tensor:
    """J(z) * direction by autograd (:12-35) -- generic path, any nn.Module."""
    z = z.requires_grad_(True)
    flat = _is_flat_decoder(decoder)

    def decode_to_image(p):
        if not flat and p.ndim == 2:
            p = p.unsqueeze(-1).unsqueeze(-1)
        img = torch.sigmoid(decoder(p))
        return img.view(img.size(0), -1)

    _, out = torch.autograd.functional.jvp(decode_to_image, (z,), (direction,))
    return out


@contextlib.contextmanager
def _strict_fp32():
    """1e-5 parity needs true fp32 convolutions/matmuls (no TF32)."""
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        yield
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b


# ------------------------------------------------------------ tensor-core GEMM for the wide decoder layer
def _lin(xin: torch.Tensor, w: torch.Tensor, normalised_next: bool) -> torch.Tensor:
    """xin @ w.  Layers with K >= 256 that feed a BatchNorm go through the hand-written tcgen05 3xTF32 GEMM
    (csrc/jvp_gemm.cu, 4x the cuBLAS SIMT SGEMM).  Its error is 1.3-1.9e-6 x sum|a||w| (the tensor core's fp32
    accumulation truncates; rounding the lo parts or adding lo.lo changes nothing, profiles/r2_gemm3x_accuracy.txt)
    against 3e-7 for SGEMM: harmless in front of a normalisation (median edge-length error 3.8e-7, kink flips 12
    instead of 9 of 15 641 on s1), but decoders without normalisation keep the library SGEMM.
    GEOCB_JVP_GEMM=cublas forces the library GEMM everywhere.  Unsupported shapes (a short last batch) fall back too."""
    if _wants_tc(w, normalised_next):
        y = dg.gemm3x(xin.contiguous(), w)
        if y is not None:
            return y
    return xin @ w


def _wants_tc(w, normalised_next):
    return w.shape[0] >= 256 and normalised_next and os.environ.get("GEOCB_JVP_GEMM", "tc") == "tc"


def _bn_fusable(ops, j, n, groups):
    nxt = ops[j] if j < len(ops) else None
    return nxt is not None and nxt[0] == "bn" and (nxt[1].training or nxt[1].running_mean is None) and n % groups == 0


# ------------------------------------------------------------ explicit forward mode
def _flatten_layers(decoder):
    """Layer list of a conv decoder, or None if it contains something we do not know."""
    out = []

    def walk(m):
        if isinstance(m, nn.Sequential):
            for c in m:
                if not walk(c):
                    return False
            return True
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d, nn.BatchNorm2d, nn.GroupNorm, nn.ReLU, nn.Identity)):
            out.append(m)
            return True
        return False

    kids = list(decoder.children())
    # SpatialDecoder.forward = deconv_layers(conv_in(z)): children order == execution order
    if not kids or not hasattr(decoder, "conv_in") or not hasattr(decoder, "deconv_layers") or len(kids) != 2:
        return None
    for k in kids:
        if not walk(k):
            return None
    return out


def _conv_pair(m, x, t):
    """(primal, tangent) through a linear layer with one library call; bias only on the primal."""
    n = x.shape[0]
    xt = torch.cat([x, t], 0)
    if isinstance(m, nn.Conv2d):
        y = F.conv2d(xt, m.weight, None, m.stride, m.padding, m.dilation, m.groups)
    else:
        y = F.conv_transpose2d(xt, m.weight, None, m.stride, m.padding, m.output_padding, m.groups, m.dilation)
    yp, yt = y[:n], y[n:]
    if m.bias is not None:
        yp = yp + m.bias.view(1, -1, 1, 1)
    return yp, yt


def _bn_pair(m, x, t, groups):
    """BatchNorm2d forward-mode.  Train mode: statistics per group of rows (= one reference batch)."""
    C = x.shape[1]
    w = m.weight.view(1, 1, C, 1, 1) if m.weight is not None else 1.0
    b = m.bias.view(1, 1, C, 1, 1) if m.bias is not None else 0.0
    if m.training or m.running_mean is None:
        xg = x.view(groups, -1, *x.shape[1:])
        tg = t.view(groups, -1, *t.shape[1:])
        mu = xg.mean(dim=(1, 3, 4), keepdim=True)
        xc = xg - mu
        var = (xc * xc).mean(dim=(1, 3, 4), keepdim=True)  # biased, as F.batch_norm normalises
        r = torch.rsqrt(var + m.eps)
        xh = xc * r
        tc = tg - tg.mean(dim=(1, 3, 4), keepdim=True)
        th = r * (tc - xh * (xh * tg).mean(dim=(1, 3, 4), keepdim=True))
        return (xh * w + b).view_as(x), (th * w).view_as(t)
    r = torch.rsqrt(m.running_var + m.eps).view(1, C, 1, 1)
    w4 = m.weight.view(1, C, 1, 1) if m.weight is not None else 1.0
    b4 = m.bias.view(1, C, 1, 1) if m.bias is not None else 0.0
    return (x - m.running_mean.view(1, C, 1, 1)) * r * w4 + b4, t * r * w4


def _gn_pair(m, x, t):
    """GroupNorm forward-mode (per-sample statistics)."""
    n, C = x.shape[:2]
    xg = x.reshape(n, m.num_groups, -1)
    tg = t.reshape(n, m.num_groups, -1)
    mu = xg.mean(-1, keepdim=True)
    xc = xg - mu
    r = torch.rsqrt((xc * xc).mean(-1, keepdim=True) + m.eps)
    xh = xc * r
    th = r * (tg - tg.mean(-1, keepdim=True) - xh * (xh * tg).mean(-1, keepdim=True))
    xh, th = xh.view_as(x), th.view_as(t)
    if m.affine:
        w, b = m.weight.view(1, C, 1, 1), m.bias.view(1, C, 1, 1)
        return xh * w + b, th * w
    return xh, th


def _flat_decoder_layers(decoder):
    """Layer list of the reference's flat-latent decoder (src/models/vae.py:53-85: fc -> view(C,4,4) -> deconv1 ->
    deconv2 -> output_layer), or None for anything else."""
    need = ("fc", "deconv1", "deconv2", "output_layer")
    if not all(hasattr(decoder, a) for a in need) or [n for n, _ in decoder.named_children()] != list(need):
        return None
    if not isinstance(decoder.fc, nn.Linear) or not isinstance(decoder.output_layer, nn.ConvTranspose2d):
        return None
    out = [decoder.fc, ("view", (-1, 4, 4))]
    for blk in (decoder.deconv1, decoder.deconv2):
        for m in blk:
            if not isinstance(m, (nn.ConvTranspose2d, nn.Conv2d, nn.BatchNorm2d, nn.GroupNorm, nn.ReLU, nn.Identity)):
                return None
            out.append(m)
    out.append(decoder.output_layer)
    return out


def _forward_mode_modules(layers, z, dz, groups):
    """Tangent of sigmoid(decoder(z)) along dz through the decoder's own layers in explicit forward mode: primal and
    tangent share every linear / convolution call (stacked along the batch), norms and ReLU propagate the tangent in
    closed form.  Used for the flat-latent decoder, whose dense unrolling would be ~10x the convolution flops."""
    n = z.shape[0]
    x, t = z, dz
    for m in layers:
        if isinstance(m, tuple):  # ("view", (C, H, W))
            x, t = x.view(n, *m[1]), t.view(n, *m[1])
        elif isinstance(m, nn.Linear):
            y = torch.cat([x, t], 0) @ m.weight.t()
            x, t = y[:n], y[n:]
            if m.bias is not None:
                x = x + m.bias
        elif isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            x, t = _conv_pair(m, x, t)
        elif isinstance(m, nn.BatchNorm2d):
            x, t = _bn_pair(m, x, t, groups)
        elif isinstance(m, nn.GroupNorm):
            x, t = _gn_pair(m, x, t)
        elif isinstance(m, nn.ReLU):
            t = t * (x > 0)
            x = torch.relu(x)
    s = torch.sigmoid(x)  # always applied (:29)
    return (s * (1 - s) * t).reshape(n, -1)


def _unique_taps(m):
    """True when every (input position, output position) pair is linked by at most one kernel tap, i.e. the dense
    matrix entries are single weights (always the case for dilation 1)."""
    return tuple(m.dilation) == (1, 1)


def _conv_matrix(m, shape):
    """Dense (C*H*W, C'*H'*W') matrix of the linear map a Conv2d / ConvTranspose2d applies to a (C,H,W) input, built
    by index arithmetic: every entry is exactly one kernel weight or 0 (no cuDNN call, whose first use costs each
    process ~1 s of library initialisation).  Returns (A, out_shape) or None for grouped convolutions."""
    if m.groups != 1 or any(isinstance(p, str) for p in (m.padding if isinstance(m.padding, tuple) else (m.padding,))):
        return None
    C, H, W = shape
    w = m.weight.detach().cpu()  # a few hundred small slice assignments: host tensors, one upload at the end
    (sh, sw), (ph, pw), (dh, dw) = m.stride, m.padding, m.dilation
    kh_n, kw_n = w.shape[2], w.shape[3]
    transposed = isinstance(m, nn.ConvTranspose2d)
    if transposed:
        oph, opw = m.output_padding
        Co = w.shape[1]
        Ho = (H - 1) * sh - 2 * ph + dh * (kh_n - 1) + oph + 1
        Wo = (W - 1) * sw - 2 * pw + dw * (kw_n - 1) + opw + 1
    else:
        Co = w.shape[0]
        Ho = (H + 2 * ph - dh * (kh_n - 1) - 1) // sh + 1
        Wo = (W + 2 * pw - dw * (kw_n - 1) - 1) // sw + 1
    A = torch.zeros(C, H, W, Co, Ho, Wo, device=w.device, dtype=w.dtype)
    for kh in range(kh_n):
        for kw in range(kw_n):
            if transposed:  # out[oh,ow] += in[ih,iw] * w[ci,co,kh,kw],  oh = ih*s - p + kh*d
                ih = torch.arange(H)
                iw = torch.arange(W)
                oh, ow = ih * sh - ph + kh * dh, iw * sw - pw + kw * dw
                wk = w[:, :, kh, kw]  # (C, Co)
            else:           # out[oh,ow] += in[oh*s - p + kh*d, ...] * w[co,ci,kh,kw]
                oh = torch.arange(Ho)
                ow = torch.arange(Wo)
                ih, iw = oh * sh - ph + kh * dh, ow * sw - pw + kw * dw
                wk = w[:, :, kh, kw].t()  # (C, Co)
            vh = (ih >= 0) & (ih < H) & (oh >= 0) & (oh < Ho)
            vw = (iw >= 0) & (iw < W) & (ow >= 0) & (ow < Wo)
            for a, b in zip(ih[vh].tolist(), oh[vh].tolist()):
                for c, d in zip(iw[vw].tolist(), ow[vw].tolist()):
                    A[:, a, c, :, b, d] = wk  # (dilation 1: one tap per (input position, output position) pair)
    return A.reshape(C * H * W, Co * Ho * Wo).contiguous(), (Co, Ho, Wo)


def _unrolled_ops(layers, latent_dim):
    """Decoder on a 1x1 latent grid as dense operators: every conv / transposed conv becomes the matrix of
    the linear map it applies to its (C,H,W) input (obtained by pushing an identity basis through the layer
    itself, so each entry is exactly one kernel weight or 0), which turns the stack into fp32 GEMMs."""
    ops, shape = [], (latent_dim, 1, 1)
    for m in layers:
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            k = shape[0] * shape[1] * shape[2]
            direct = _conv_matrix(m, shape) if _unique_taps(m) else None
            if direct is not None:
                A, oshape = direct
            else:
                eye = torch.eye(k, device=m.weight.device, dtype=m.weight.dtype).view(k, *shape)
                if isinstance(m, nn.Conv2d):
                    out = F.conv2d(eye, m.weight, None, m.stride, m.padding, m.dilation, m.groups)
                else:
                    out = F.conv_transpose2d(eye, m.weight, None, m.stride, m.padding, m.output_padding, m.groups,
                                             m.dilation)
                oshape = tuple(out.shape[1:])
                A = out.reshape(k, -1).contiguous()
            A = A.cpu()
            bias = None if m.bias is None else m.bias.detach().cpu().view(-1, 1, 1).expand(oshape).reshape(-1).contiguous()
            if ops and ops[-1][0] == "lin":
                # two linear maps in a row (conv_in -> first ConvT): compose once, in fp64, round once
                _, A0, b0, _ = ops.pop()
                A64 = A0.double() @ A.double()
                if b0 is not None:
                    b64 = b0.double() @ A.double() + (0 if bias is None else bias.double())
                    bias = b64.float()
                A = A64.float().contiguous()
            ops.append(("lin", A, bias, oshape))
            shape = oshape
        elif isinstance(m, nn.BatchNorm2d):
            ops.append(("bn", m, shape))
        elif isinstance(m, nn.GroupNorm):
            ops.append(("gn", m, shape))
        elif isinstance(m, nn.ReLU):
            ops.append(("relu",))
    dev = next((p.device for m in layers for p in m.parameters()), torch.device("cpu"))
    return [("lin", o[1].to(dev).contiguous(), None if o[2] is None else o[2].to(dev), o[3]) if o[0] == "lin" else o
            for o in ops]


def _forward_mode(ops, z, dz, groups):
    """Tangent of sigmoid(decoder(z)) along dz for rows z (R,C); returns (R,P).
    lin -> BatchNorm(train) -> ReLU runs as one GEMM + one fused in-place kernel (`geocb_bn_relu_jvp`)."""
    x, t = z, dz
    n = z.shape[0]
    pending_bias = None  # bias of the last linear map, folded into the fused BN kernel when one follows
    stacked = None       # buffer whose halves x and t currently are (the fused BN kernel works in place on it)
    in_image = False     # the BN kernel has written the next GEMM's operand image instead of x / t
    i = 0
    while i < len(ops):
        op = ops[i]
        if op[0] == "lin":
            fusable = _bn_fusable(ops, i + 1, n, groups)
            if i == 0 and fusable and op[1].shape[0] in (16, 32) and os.environ.get("GEOCB_JVP_FUSE_IN", "1") != "0":
                # first layer (latent -> first feature map) + BatchNorm + ReLU + operand image of the wide GEMM behind
                # them in ONE kernel: the pre-activation matrix is never materialised
                m, shape = ops[1][1], ops[1][2]
                relu_next = len(ops) > 2 and ops[2][0] == "relu"
                j = 3 if relu_next else 2
                if (j < len(ops) and ops[j][0] == "lin" and _wants_tc(ops[j][1], _bn_fusable(ops, j + 1, n, groups))
                        and n % 128 == 0 and ops[j][1].shape[0] == shape[0] * shape[1] * shape[2]
                        and dg.gemm3x_supported(2 * n, ops[j][1].shape[0], ops[j][1].shape[1])):
                    image = (dg.gemm3x_image_buffer(2 * n, ops[j][1].shape[0], x.device), n)
                    if dg.lin_bn_relu_jvp_to_image(x.contiguous(), t.contiguous(), op[1], shape[0], shape[1] * shape[2],
                                                   n // groups, op[2], m.weight, m.bias, m.eps, relu_next, image):
                        in_image = True
                        i = j
                        continue
            if (i == len(ops) - 1 and not in_image and os.environ.get("GEOCB_JVP_FUSE_OUT", "1") != "0"
                    and x.shape == t.shape):
                # last layer + sigmoid + its tangent in one kernel (narrow outputs: the library SGEMM is 2.4x off the
                # memory time there, and it would be the only library GEMM left on the path)
                j = dg.lin_sigmoid_jvp(x.contiguous(), t.contiguous(), op[1], op[2])
                if j is not None:
                    return j
            if in_image:
                y = dg.gemm3x(None, op[1], M=2 * n)
                in_image = False
            else:
                xin = stacked if stacked is not None else torch.cat([x, t], 0)  # primal and tangent share one GEMM
                y = _lin(xin, op[1], fusable)
            x, t = y[:n], y[n:]
            stacked = y
            pending_bias = op[2]
            if not fusable and pending_bias is not None:
                x = x + pending_bias
                pending_bias = None
                stacked = None
        elif op[0] == "bn":
            m, shape = op[1], op[2]
            relu_next = i + 1 < len(ops) and ops[i + 1][0] == "relu"
            done = False
            if (m.training or m.running_mean is None) and n % groups == 0:
                # when the layer after this BN(+ReLU) is a wide linear map bound for the tcgen05 GEMM, let the BN
                # kernel write that GEMM's tf32 hi | lo operand image directly (no row-major round trip, no split pass)
                j = i + 2 if relu_next else i + 1
                image = None
                if (j < len(ops) and ops[j][0] == "lin" and _wants_tc(ops[j][1], _bn_fusable(ops, j + 1, n, groups))
                        and n % 128 == 0 and ops[j][1].shape[0] == shape[0] * shape[1] * shape[2]
                        and dg.gemm3x_supported(2 * n, ops[j][1].shape[0], ops[j][1].shape[1])):
                    image = (dg.gemm3x_image_buffer(2 * n, ops[j][1].shape[0], x.device), n)
                done = dg.bn_relu_jvp(x, t, shape[0], shape[1] * shape[2], n // groups, pending_bias, m.weight, m.bias,
                                      m.eps, relu_next, image=image)
                in_image = done and image is not None
            if done:
                pending_bias = None
                if relu_next:
                    i += 1
            else:
                if pending_bias is not None:
                    x = x + pending_bias
                    pending_bias = None
                x4, t4 = _bn_pair(m, x.reshape(n, *shape), t.reshape(n, *shape), groups)
                x, t = x4.reshape(n, -1), t4.reshape(n, -1)
                stacked = None
        elif op[0] == "gn":
            x4, t4 = _gn_pair(op[1], x.reshape(n, *op[2]), t.reshape(n, *op[2]))
            x, t = x4.reshape(n, -1), t4.reshape(n, -1)
            stacked = None
        else:
            t = t * (x > 0)
            x = torch.relu(x)
            stacked = None
        i += 1
    s = torch.sigmoid(x)  # always applied (:29)
    return s * (1 - s) * t


def decoder_needs_library_gemm(decoder, latent_dim) -> bool:
    """Cheap structural guess (shapes only, no tensor work): will the edge-length stage call a library GEMM for this
    decoder?  It does not when the first layer fuses into the BatchNorm kernel (latent_dim 16 / 32), the middle layers
    run on the tcgen05 GEMM and the output patch of a latent cell has 16 values (geocb_lin_sigmoid_jvp).  Used by the
    pipeline to decide whether creating the library handle in the kNN kernel's shadow is worth 0.1 s of host time."""
    if _is_flat_decoder(decoder):
        return True
    layers = _flatten_layers(decoder)
    if layers is None or latent_dim not in (16, 32):
        return True
    shape = (latent_dim, 1, 1)
    for m in layers:
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            if m.groups != 1 or isinstance(m.padding, str):
                return True
            (sh, sw), (ph, pw), (dh, dw) = m.stride, m.padding, m.dilation
            kh, kw = m.kernel_size
            if isinstance(m, nn.ConvTranspose2d):
                oph, opw = m.output_padding
                shape = (m.out_channels, (shape[1] - 1) * sh - 2 * ph + dh * (kh - 1) + oph + 1,
                         (shape[2] - 1) * sw - 2 * pw + dw * (kw - 1) + opw + 1)
            else:
                shape = (m.out_channels, (shape[1] + 2 * ph - dh * (kh - 1) - 1) // sh + 1,
                         (shape[2] + 2 * pw - dw * (kw - 1) - 1) // sw + 1)
    return shape[0] * shape[1] * shape[2] != 16


def decoder_plan(decoder, latent_dim):
    """(dense-operator list | None, flat-decoder layer list | None) for `decoder`, cached on the module and rebuilt
    when a parameter or the train/eval mode changes.  Building it costs tens of milliseconds of host work (and the
    first GEMM creates the cuBLAS handle), which the pipeline overlaps with the kNN kernel (warm_up_decoder)."""
    key = (latent_dim, tuple((p.data_ptr(), p._version) for p in decoder.parameters()),
           tuple(m.training for m in decoder.modules()))
    cached = getattr(decoder, "_geocb_plan", None)
    if cached is not None and cached[0] == key:
        return cached[1], cached[2]
    flat_layers = _flat_decoder_layers(decoder) if _is_flat_decoder(decoder) else None
    layers = None if _is_flat_decoder(decoder) else _flatten_layers(decoder)
    if layers is not None:
        with _strict_fp32():
            layers = _unrolled_ops(layers, latent_dim)
    object.__setattr__(decoder, "_geocb_plan", (key, layers, flat_layers))
    return layers, flat_layers


@torch.no_grad()
def warm_up_decoder(decoder, latent_dim, batch_size=512):
    """Build the plan and push one tiny batch through it (current stream): creates library handles, weight images and
    the allocator's first blocks off the critical path.  Results are discarded; BatchNorm buffers are not touched
    (the explicit forward mode never calls the modules)."""
    layers, flat = decoder_plan(decoder, latent_dim)
    if layers is None:
        return
    dev = next(decoder.parameters()).device
    z = torch.zeros((2 * batch_size, latent_dim), device=dev)
    with _strict_fp32():
        _forward_mode(layers, z, z, 2)


@torch.no_grad()
def edge_lengths_device(decoder, z_flat, ei, ej, batch_size=512, pos_ij=None, pos_ji=None, csr_data=None,
                        e_lo=0, e_hi=None, groups_per_call=64):
    """Edge lengths for edges [e_lo,e_hi) of the (ei,ej) list, all tensors on the decoder's CUDA device.

    Batches are the reference's: consecutive `batch_size` edges of the GLOBAL list, so e_lo must be a
    multiple of batch_size (multi-GPU sharding keeps whole batches).  Returns len f32[e_hi-e_lo]; when
    csr_data/pos_* are given the lengths are also scattered to both CSR directions.
    """
    E = ei.numel()
    e_hi = E if e_hi is None else e_hi
    assert e_lo % batch_size == 0
    dev = z_flat.device
    out = torch.empty(max(e_hi - e_lo, 0), dtype=torch.float32, device=dev)
    layers, flat_layers = decoder_plan(decoder, z_flat.shape[1])
    if flat_layers is not None:
        groups_per_call = min(groups_per_call, 8)  # full-resolution feature maps: bound the activation memory
    with _strict_fp32():
        lo = e_lo
        while lo < e_hi:
            full = (e_hi - lo) // batch_size
            if (layers is not None or flat_layers is not None) and full >= 1:
                g = min(full, groups_per_call)
                hi = lo + g * batch_size
            else:
                g, hi = 1, min(lo + batch_size, e_hi)
            a = z_flat[ei[lo:hi].long()]
            b = z_flat[ej[lo:hi].long()]
            d = b - a
            if layers is not None:
                # start- and end-point JVPs are separate decoder calls in the reference => separate BN groups
                j = _forward_mode(layers, torch.cat([a, b], 0), torch.cat([d, d], 0), 2 * g)
                ja, jb = j[: hi - lo].contiguous(), j[hi - lo:].contiguous()
            elif flat_layers is not None:
                j = _forward_mode_modules(flat_layers, torch.cat([a, b], 0), torch.cat([d, d], 0), 2 * g)
                ja, jb = j[: hi - lo].contiguous(), j[hi - lo:].contiguous()
            else:
                with torch.enable_grad():
                    ja = _compute_jacobian_vector_product(decoder, a.clone(), d).contiguous()
                    jb = _compute_jacobian_vector_product(decoder, b.clone(), d).contiguous()
            dg.edge_len_scatter(ja.float(), jb.float(), lo - e_lo, None, None, out, None)
            if csr_data is not None:
                dg.edge_len_scatter(ja.float(), jb.float(), lo, pos_ij, pos_ji, None, csr_data)
            lo = hi
    return out


@torch.no_grad()
def edge_lengths_riemannian(decoder, z_start: torch.Tensor, z_end: torch.Tensor, batch_size: int = 512) -> torch.Tensor:
    """Calculate Riemannian edge lengths between latent points (:37-66).

    Returns float32 (E,) on the decoder's device.  A CPU-resident decoder is copied to the GPU for the
    computation (there is no CPU path) and the result is moved back to its device."""
    assert z_start.shape == z_end.shape, "Start and end points must have same shape"  # :40
    home = next(decoder.parameters()).device  # :42
    dev = _lib.require_cuda()
    dec = decoder if home.type == "cuda" else copy.deepcopy(decoder).to(dev)
    dev = next(dec.parameters()).device
    E = z_start.size(0)
    if E == 0:
        return torch.empty(0, dtype=torch.float32, device=home)
    # stack start/end rows as a 2E-row table so the shared device routine can index it
    table = torch.cat([z_start.to(dev, torch.float32), z_end.to(dev, torch.float32)], 0).contiguous()
    ei = torch.arange(E, device=dev, dtype=torch.int32)
    ej = ei + E
    out = edge_lengths_device(dec, table, ei, ej, batch_size=batch_size)
    return out.to(home).to(torch.float32)
