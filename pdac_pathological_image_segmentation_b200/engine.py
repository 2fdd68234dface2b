"""Execution engine of the drop-in modules: a tape of hand-written-kernel launches (ops.py -> libpdac_b200.so) with
hand-derived backward steps. No torch.nn.functional, no cuDNN, no autograd inside: torch provides device memory
(caching allocator) and the current stream only. The modules in resunet.py / smp.py write their forward as straight
calls of the layer functions below; each call launches its kernels and, when gradients are wanted, records a closure
that launches the matching backward kernels. Gradients of parameters are accumulated into a flat fp32 arena whose
slots have the parameters' own (state_dict) layouts, so they can be handed to autograd / DDP / Adam as they are.

Activations: NHWC bf16 torch tensors (N, H, W, C) with C a multiple of 16 (logit heads are zero-padded to 16)."""
from __future__ import annotations

import torch

from . import ops
from .containers import BatchNorm2d, Conv2d, ConvTranspose2d

F32 = torch.float32


def pad16(c):
    return (c + 15) // 16 * 16


# ------------------------------------------------------------------------------------------------ packed weights
class PackCache:
    """bf16 K-major copies of the fp32 state_dict weights in the layouts the tcgen05 kernels read (ops.pack_*),
    refreshed when a parameter's storage or version counter changes (optimizer.step / load_state_dict / .to()).
    The trainer, which updates parameters through raw pointers, calls refresh() after its optimizer step: every cached
    copy is re-packed IN PLACE (same buffers, so a captured CUDA graph stays valid), the conv weights by one batched
    launch."""

    def __init__(self):
        self._c = {}

    def invalidate(self):
        """Forget every packed copy (parameters changed behind torch's version counters, e.g. a checkpoint load)."""
        self._c.clear()

    def refresh(self):
        """Re-pack every cached copy from the current parameter values into its existing buffers."""
        recipes = []
        for key, (sig, val, redo) in self._c.items():
            r = redo()
            if r is not None:
                recipes.append(r)
        ops.pack_conv_weights_batched(recipes)

    def _get(self, key, params, build, redo):
        sig = tuple((p.data_ptr(), p._version) for p in params if p is not None)
        hit = self._c.get(key)
        if hit is not None:
            if hit[0] == sig:
                return hit[1]
            if tuple(s[0] for s in hit[0]) == tuple(s[0] for s in sig):
                # same storage, new values (optimizer.step / load_state_dict): re-pack into the existing buffers, which
                # a captured CUDA graph may have baked in
                r = hit[2]()
                if r is not None:
                    ops.pack_conv_weights_batched([r])
                self._c[key] = (sig, hit[1], hit[2])
                return hit[1]
        val = build()
        self._c[key] = (sig, val, lambda: redo(val))
        return val

    def conv(self, m: Conv2d, want_dgrad=True):
        """-> (wp_f [R*S*wrows][Cin], wp_t [R*S*wrows_t][pad16(Cout)] or None, bias padded to pad16(Cout) or None)."""
        cpad = pad16(m.cout)

        def build():
            w = m.weight.detach()
            wp_f, wp_t = ops.pack_conv_weight(w.contiguous(), want_fwd=True, want_dgrad=want_dgrad, stride=m.stride,
                                              pad=m.padding, cout_pad=cpad)
            bias = None
            if m.bias is not None:
                if m.cout == cpad:
                    bias = m.bias.detach()
                elif getattr(m.bias, "_pdac_padded_slot", False):
                    # the trainer keeps parameters in zero-padded >= 16-float arena slots: read the pad in place
                    bias = torch.as_strided(m.bias.detach(), (cpad,), (1,))
                else:
                    bias = torch.zeros(cpad, dtype=F32, device=w.device)
                    bias[:m.cout].copy_(m.bias.detach())
            return wp_f, wp_t, bias

        def redo(val):
            wp_f, wp_t, bias = val
            if bias is not None and bias.data_ptr() != m.bias.data_ptr():
                bias[:m.cout].copy_(m.bias.detach())          # private zero-padded copy (module path only)
            return ops.conv_pack_recipe(m.weight.detach(), wp_f, wp_t, stride=m.stride, pad=m.padding, cout_pad=cpad)
        return self._get((id(m), want_dgrad), (m.weight, m.bias), build, redo)

    def bn_eval(self, bn: BatchNorm2d, compute):
        """Eval-mode BatchNorm as (scale / shift, mean / invstd) tensors, computed once per parameter version instead of by
        one bn_prepare launch per layer and forward (30 launches, 0.2 of the 5.5 ms of a whole-slide inference batch).
        compute(ss, mi) fills the two [2, C] tensors. Training-mode statistics updates drop the entry (drop_bn)."""
        def build():
            c = bn.num_features
            dev = bn.weight.device
            val = (torch.empty((2, c), dtype=F32, device=dev), torch.empty((2, c), dtype=F32, device=dev))
            compute(*val)
            return val

        def redo(val):
            compute(*val)
            return None
        return self._get(("bn", id(bn)), (bn.weight, bn.bias, bn.running_mean, bn.running_var), build, redo)

    def drop_bn(self, bn: BatchNorm2d):
        self._c.pop(("bn", id(bn)), None)

    def convt(self, m: ConvTranspose2d):
        def build():
            wp_f, wp_t = ops.pack_convt_weight(m.weight.detach().contiguous())
            return wp_f, wp_t, m.bias.detach()

        def redo(val):
            ops.pack_convt_weight(m.weight.detach(), wp_f=val[0], wp_t=val[1])
            return None
        return self._get((id(m), "t"), (m.weight, m.bias), build, redo)

    def stem(self, m: Conv2d):
        def redo(val):
            ops.pack_stem_weight(m.weight.detach(), wp=val)
            return None
        return self._get((id(m), "stem"), (m.weight,), lambda: ops.pack_stem_weight(m.weight.detach().contiguous()), redo)


# ------------------------------------------------------------------------------------------------ gradient arena
class GradArena:
    """One flat fp32 buffer with a slot per parameter (slot = the parameter's own layout, 256-byte aligned, at least
    16 floats so channel-padded bias reductions stay inside their slot)."""

    def __init__(self, params, device):
        self.offsets = {}
        off = 0
        for p in params:
            n = max(p.numel(), 16)
            self.offsets[id(p)] = (off, p.numel(), tuple(p.shape))
            off += (n + 63) // 64 * 64
        self.size = off
        self.flat = torch.zeros(max(off, 64), dtype=F32, device=device)
        self._views = {}

    def zero(self):
        self.flat.zero_()

    def slot(self, p, padded=False):
        """fp32 view with the parameter's shape (or, padded=True, the raw >=16-float slot as 1-D)."""
        off, n, shape = self.offsets[id(p)]
        if padded:
            return self.flat[off:off + max(n, 16)]
        v = self._views.get(id(p))
        if v is None:
            v = self.flat[off:off + n].view(shape)
            self._views[id(p)] = v
        return v


# ------------------------------------------------------------------------------------------------ tape
class Tape:
    def __init__(self, packs: PackCache, training: bool, arena: GradArena | None, step_dev=None, base_seed=0):
        """step_dev / base_seed: the trainer's device-side step counter and a per-trainer seed; with them the random
        layers (Dropout) draw no host numbers per step, so the step can be replayed from a CUDA graph."""
        self.packs = packs
        self.training = training
        self.arena = arena
        self.step_dev = step_dev
        self.base_seed = int(base_seed)
        self.n_random = 0
        self.enabled = arena is not None
        self.steps = []
        self.grads = {}
        self.nograd = set()
        self.host_rng = False
        # nearest x2 upsamplings: ids of their outputs (forward), and of those whose pending gradient is already the
        # half-resolution 2x2 sum (written by the consuming conv's fused data-gradient kernel)
        self.up2_src = set()
        self.pooled = set()
        # BatchNorm partial-sum tables of the conv epilogues: the trainer hands over ONE zeroed buffer per step (sized
        # from the previous step's demand) instead of ~30 separate fills; without it every table is its own torch.zeros
        self.stats_pool = None
        self.stats_used = 0

    def new_stats(self, c, device):
        """Zeroed [STAT_ROWS, 2, c] table (ops.new_stats), carved from the step's pool while it lasts."""
        n = ops.STAT_ROWS * 2 * c
        off = self.stats_used
        self.stats_used += n
        pool = self.stats_pool
        if pool is not None and off + n <= pool.numel() and pool.device == torch.device(device):
            return pool[off:off + n].view(ops.STAT_ROWS, 2, c)
        return ops.new_stats(c, device)

    def record(self, fn, params=()):
        """fn: the layer's backward step; params: the parameters whose gradient slots that step completes (the
        data-parallel reducer launches a bucket's allreduce as soon as all of its parameters are complete)."""
        if self.enabled:
            self.steps.append((fn, tuple(p for p in params if p is not None)))

    def no_grad_for(self, t):
        self.nograd.add(id(t))

    def wants(self, t):
        return id(t) not in self.nograd

    def put(self, t, g):
        if id(t) in self.nograd:
            return
        k = id(t)
        old = self.grads.get(k)
        if old is None:
            self.grads[k] = g
        elif k in self.pooled:        # the pending gradient is already pooled: pool this one into it
            ops.upsample_nearest2x_bwd(g, old, out=old)
        else:
            ops.add(old, g, out=old)

    def put_pooled(self, t, g):
        """g: gradient of the x2-upsampled tensor t, already summed over the 2x2 replicas (half resolution)."""
        if id(t) in self.nograd:
            return
        k = id(t)
        old = self.grads.get(k)
        if old is None:
            self.grads[k] = g
        elif k in self.pooled:
            ops.add(old, g, out=old)
        else:                         # a full-resolution gradient is pending: pool it into g
            self.grads[k] = ops.upsample_nearest2x_bwd(old, g, out=g)
        self.pooled.add(k)

    def pop(self, t):
        return self.grads.pop(id(t), None)

    def pop_up2(self, t):
        """-> (pending gradient of the x2-upsampled tensor t or None, whether it is already pooled)."""
        k = id(t)
        pooled = k in self.pooled
        self.pooled.discard(k)
        return self.grads.pop(k, None), pooled

    def backward(self, on_params_done=None):
        for fn, params in reversed(self.steps):
            fn()
            if on_params_done is not None and params:
                on_params_done(params)
        self.steps = []
        self.grads = {}
        self.pooled = set()


# ------------------------------------------------------------------------------------------------ layer functions
def _conv_backward(tape, m: Conv2d, srcs, dy, wp_t):
    """wgrad into the arena slot (state_dict layout) and dgrad to the sources; dy has pad16(Cout) channels."""
    k = m.k
    ops.conv2d_wgrad(srcs, dy, k, k, m.stride, m.padding, m.dilation, out=tape.arena.slot(m.weight), oihw=True,
                     cout_valid=m.cout)
    need = [tape.wants(s) for s in srcs]
    if not any(need):
        return
    h, w = srcs[0].shape[1], srcs[0].shape[2]
    s0 = srcs[0]
    if (id(s0) in tape.up2_src and all(need) and len(srcs) <= 2 and m.stride == 1 and id(s0) not in tape.grads
            and 2 * m.padding == (k - 1) * m.dilation):
        # the first source is a nearest x2 upsampling: its gradient leaves the data-gradient kernel already summed over
        # the 2x2 replicas, at half resolution (no full-resolution round trip through upsample_nearest2x_bwd)
        c_rest = srcs[1].shape[3] if len(srcs) == 2 else 0
        res = ops.conv2d_dgrad_pool2(dy, wp_t, s0.shape[3], c_rest, k, k, m.padding, m.dilation, c_real=m.cout)
        if res is not None:
            tape.put_pooled(s0, res[0])
            if c_rest:
                tape.put(srcs[1], res[1])
            return
    if len(srcs) == 1:
        old = tape.pop(srcs[0])
        dx = ops.conv2d_dgrad(dy, wp_t, srcs[0].shape[3], (h, w), k, k, m.stride, m.padding, m.dilation, addend=old,
                              out=old, c_real=m.cout)
        tape.put(srcs[0], dx)
    else:
        ctot = sum(s.shape[3] for s in srcs)
        dx = ops.conv2d_dgrad(dy, wp_t, ctot, (h, w), k, k, m.stride, m.padding, m.dilation, c_real=m.cout)
        c0 = 0
        for s in srcs:
            tape.put(s, dx[..., c0:c0 + s.shape[3]])
            c0 += s.shape[3]


def conv_bias_act(tape: Tape, m: Conv2d, srcs, relu):
    """[relu](conv(cat(srcs)) + bias) with everything fused in the GEMM epilogue (models/resunet.py:38,42,46,48)."""
    wp_f, wp_t, bias = tape.packs.conv(m)
    k = m.k
    y = ops.conv2d_fwd(srcs, wp_f, pad16(m.cout), k, k, m.stride, m.padding, m.dilation, relu=relu, bias=bias,
                       c_real=m.cout)

    def bwd():
        dy = tape.pop(y)
        if dy is None:
            return
        if relu or m.bias is not None:
            dbias = tape.arena.slot(m.bias, padded=True)[:dy.shape[3]] if m.bias is not None else None
            ops.relu_bwd_bias(dy, y if relu else None, dy if relu else None, dbias)
        _conv_backward(tape, m, srcs, dy, wp_t)
    tape.record(bwd, (m.weight, m.bias))
    return y


class _BNState:
    __slots__ = ("ss", "mi", "count")


def _bn_prepare(tape, bn: BatchNorm2d, stats, count, device):
    st = _BNState()
    c = bn.num_features
    if not tape.training:      # running statistics: a pure function of the module's parameters and buffers -> cached
        def compute(ss, mi):
            ops.bn_prepare(None, 1, bn.weight.detach(), bn.bias.detach(), bn.eps, bn.momentum, False,
                           bn.running_mean, bn.running_var, ss, mi, num_batches_tracked=None)
        st.ss, st.mi = tape.packs.bn_eval(bn, compute)
        st.count = count
        return st
    tape.packs.drop_bn(bn)     # this call updates the running statistics through raw pointers
    st.ss = torch.empty((2, c), dtype=F32, device=device)
    st.mi = torch.empty((2, c), dtype=F32, device=device)
    st.count = count
    if tape.training and count <= 1:
        raise ValueError(f"Expected more than 1 value per channel when training, got input size [{count}, {c}]")
    ops.bn_prepare(stats, count, bn.weight.detach(), bn.bias.detach(), bn.eps, bn.momentum, tape.training,
                   bn.running_mean, bn.running_var, st.ss, st.mi,
                   num_batches_tracked=bn.num_batches_tracked if tape.training else None)
    return st


def conv_bn_act(tape: Tape, m: Conv2d, bn: BatchNorm2d, srcs, relu=True, residual=None, out=None, up2=False):
    """[relu](BN(conv(cat(srcs))) + residual): conv with per-channel sum / sum-of-squares in its epilogue, a finalize
    kernel, then one fused normalise(+residual)(+ReLU) pass (torchvision resnet.py:92-103; smp Conv2dReLU).
    up2 (training path): returns the nearest-neighbour x2 upsampling of the result, written by the normalise pass itself
    (smp UnetDecoder: F.interpolate(x, scale_factor=2, mode="nearest") in front of the next block)."""
    wp_f, wp_t, _ = tape.packs.conv(m)
    k = m.k
    dev = srcs[0].device
    if not tape.training and not tape.enabled:
        # inference: BatchNorm is a per-channel affine map -> folded (with the residual and the ReLU) into the conv epilogue
        st = _bn_prepare(tape, bn, None, 1, dev)
        if up2 and residual is None and out is None and m.stride == 1:
            # the next decoder block's upsampling written by this conv's epilogue (every pixel stored four times)
            y2 = ops.conv2d_fwd_affine_up2(srcs, wp_f, m.cout, k, k, m.padding, m.dilation, relu, st.ss[0], st.ss[1])
            if y2 is not None:
                return y2
        y = ops.conv2d_fwd(srcs, wp_f, m.cout, k, k, m.stride, m.padding, m.dilation, relu=relu, bias=st.ss[1],
                           addend=residual, scale=st.ss[0], out=out)
        return ops.upsample_nearest2x_fwd(y) if up2 else y
    stats = tape.new_stats(m.cout, dev) if tape.training else None
    x = ops.conv2d_fwd(srcs, wp_f, m.cout, k, k, m.stride, m.padding, m.dilation, stats=stats)
    st = _bn_prepare(tape, bn, stats, x.shape[0] * x.shape[1] * x.shape[2], dev)
    keep_x = tape.enabled
    if out is not None:        # write the activation straight into a channel slice of a concat buffer
        assert not up2
        y = ops.scale_shift_act(x, st.ss, residual, relu=relu, out=out)
    elif up2:
        y = ops.scale_shift_act(x, st.ss, residual, relu=relu, up2=True)
    else:
        y = ops.scale_shift_act(x, st.ss, residual, relu=relu, out=None if keep_x else x)
    training = tape.training

    if up2 and tape.enabled:
        tape.up2_src.add(id(y))

    def bwd():
        dy, pooled = tape.pop_up2(y) if up2 else (tape.pop(y), False)
        if dy is None:
            return
        if up2 and not pooled:  # gradient of the fused upsampling: sum of the four replicas
            dy = ops.upsample_nearest2x_bwd(dy)
        sums = ops.new_bwd_sums(m.cout, dev)
        # without a residual the ReLU mask is recomputed from x and the forward's scale / shift: y is not re-read
        assert not (up2 and residual is not None), "up2 is fused for residual-free layers only (mask recomputed from x)"
        ymask = y if (relu and residual is not None) else None
        ssm = st.ss if (relu and residual is None) else None
        ops.bn_bwd_reduce(dy, ymask, x, st.mi, sums, scale_shift=ssm)
        want_res = residual is not None and tape.wants(residual)
        dxr = torch.empty_like(dy) if want_res else dy
        ops.bn_bwd_apply(dy, ymask, x, st.mi, bn.weight.detach(), sums, st.count, training, dx=dxr,
                         dz_out=dy if (want_res and relu) else None, dgamma=tape.arena.slot(bn.weight),
                         dbeta=tape.arena.slot(bn.bias), scale_shift=ssm)
        if want_res:
            tape.put(residual, dy)
        _conv_backward(tape, m, srcs, dxr, wp_t)
    tape.record(bwd, (m.weight, bn.weight, bn.bias))
    return y


def stem(tape: Tape, m: Conv2d, bn: BatchNorm2d, z):
    """7x7 s2 p3 conv + BN + ReLU on the space-to-depth input z (N, H/2, W/2, 16) as a 4x4 s1 tap convolution."""
    wp = tape.packs.stem(m)
    dev = z.device
    if not tape.training and not tape.enabled:
        st = _bn_prepare(tape, bn, None, 1, dev)
        return ops.stem_conv_fwd(z, wp, m.cout, scale=st.ss[0], shift=st.ss[1], relu=True)
    stats = tape.new_stats(m.cout, dev) if tape.training else None
    x = ops.stem_conv_fwd(z, wp, m.cout, stats=stats)
    st = _bn_prepare(tape, bn, stats, x.shape[0] * x.shape[1] * x.shape[2], dev)
    y = ops.scale_shift_act(x, st.ss, None, relu=True, out=None if tape.enabled else x)
    training = tape.training

    def bwd():
        dy = tape.pop(y)
        if dy is None:
            return
        sums = ops.new_bwd_sums(m.cout, dev)
        ops.bn_bwd_reduce(dy, None, x, st.mi, sums, scale_shift=st.ss)
        ops.bn_bwd_apply(dy, None, x, st.mi, bn.weight.detach(), sums, st.count, training, dx=dy,
                         dgamma=tape.arena.slot(bn.weight), dbeta=tape.arena.slot(bn.bias), scale_shift=st.ss)
        dwp = ops.stem_conv_wgrad(z, dy)
        ops.unpack_stem_wgrad(dwp, m.cout, out=tape.arena.slot(m.weight))
    tape.record(bwd, (m.weight, bn.weight, bn.bias))
    return y


def maxpool(tape: Tape, x):
    y, idx = ops.maxpool_fwd(x, want_idx=tape.enabled)

    def bwd():
        dy = tape.pop(y)
        if dy is not None:
            old = tape.pop(x)        # e.g. the decoder's skip-connection gradient: added inside the scatter pass
            tape.put(x, ops.maxpool_bwd(dy, idx, (x.shape[1], x.shape[2]), addend=old))
    tape.record(bwd)
    return y


def convt2x2(tape: Tape, m: ConvTranspose2d, x):
    """ConvTranspose2d(k=2, s=2) + bias: four parity-class GEMMs with pixel-shuffle stores (models/resunet.py:36,40,44)."""
    wp_f, wp_t, bias = tape.packs.convt(m)
    y = ops.convt2x2_fwd(x, wp_f, m.cout, bias=bias)

    def bwd():
        dy = tape.pop(y)
        if dy is None:
            return
        ops.relu_bwd_bias(dy, None, None, tape.arena.slot(m.bias))
        ops.convt2x2_wgrad(x, dy, out=tape.arena.slot(m.weight), iohw=True)
        if tape.wants(x):
            tape.put(x, ops.convt2x2_dgrad(dy, wp_t, m.cin))
    tape.record(bwd, (m.weight, m.bias))
    return y


def nearest2x(tape: Tape, x):
    y = ops.upsample_nearest2x_fwd(x)
    if tape.enabled:
        tape.up2_src.add(id(y))

    def bwd():
        dy, pooled = tape.pop_up2(y)
        if dy is not None:
            if pooled:
                tape.put(x, dy)
            else:
                old = tape.pop(x)
                tape.put(x, ops.upsample_nearest2x_bwd(dy, old, out=old))
    tape.record(bwd)
    return y


def bilinear(tape: Tape, x, out_hw, align_corners):
    y = ops.bilinear_fwd(x, out_hw, align_corners)

    def bwd():
        dy = tape.pop(y)
        if dy is not None:
            tape.put(x, ops.bilinear_bwd(dy, (x.shape[1], x.shape[2]), align_corners))
    tape.record(bwd)
    return y


def add(tape: Tape, a, b):
    y = ops.add(a, b)

    def bwd():
        dy = tape.pop(y)
        if dy is not None:
            tape.put(a, dy)
            tape.put(b, dy.clone() if tape.wants(a) and tape.wants(b) else dy)
    tape.record(bwd)
    return y


def conv_gn_relu(tape: Tape, m: Conv2d, gn, srcs):
    """smp Conv3x3GNReLU body: relu(GroupNorm(conv(x))); per-image group statistics in fp32."""
    wp_f, wp_t, _ = tape.packs.conv(m)
    k = m.k
    x = ops.conv2d_fwd(srcs, wp_f, m.cout, k, k, m.stride, m.padding, m.dilation)
    y, stats = ops.groupnorm_fwd(x, gn.num_groups, gn.weight.detach(), gn.bias.detach(), gn.eps, relu=True)

    def bwd():
        dy = tape.pop(y)
        if dy is None:
            return
        dx = ops.groupnorm_bwd(dy, y, x, gn.num_groups, stats, gn.weight.detach(), gn.eps,
                               dgamma=tape.arena.slot(gn.weight), dbeta=tape.arena.slot(gn.bias), out=dy)
        _conv_backward(tape, m, srcs, dx, wp_t)
    tape.record(bwd, (m.weight, gn.weight, gn.bias))
    return y


def dropout(tape: Tape, x, p, per_channel):
    """nn.Dropout2d / nn.Dropout: identity in eval mode or with p == 0; in training a hash-generated mask (the seed is
    drawn from torch's CPU generator, so torch.manual_seed makes runs reproducible; it is NOT torch's Philox stream)."""
    if not tape.training or p <= 0.0:
        return x
    step = tape.step_dev
    if step is not None:
        # trainer path: the seed is (trainer seed, index of this layer call); the device step counter is mixed in by the kernel
        seed = (tape.base_seed + 0x9E3779B97F4A7C15 * (tape.n_random + 1)) & (2 ** 62 - 1)
        tape.n_random += 1
    else:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        tape.host_rng = True   # a per-step host value is baked into the launch: such a step must not be replayed from a graph
    y = ops.dropout(x, p, seed, per_channel, step=step)

    def bwd():
        dy = tape.pop(y)
        if dy is not None:
            tape.put(x, ops.dropout(dy, p, seed, per_channel, out=dy, step=step))
    tape.record(bwd)
    return y


def depthwise3x3(tape: Tape, m: Conv2d, x):
    """SeparableConv2d's depthwise half: Conv2d(C, C, 3, padding=d, dilation=d, groups=C, bias=False)."""
    w = m.weight.detach()
    y = ops.dwconv3x3(x, w, m.dilation)

    def bwd():
        dy = tape.pop(y)
        if dy is None:
            return
        ops.dwconv3x3_wgrad(x, dy, m.dilation, tape.arena.slot(m.weight))
        if tape.wants(x):
            old = tape.pop(x)
            dx = ops.dwconv3x3(dy, w, m.dilation, flip=True)
            if old is not None:
                ops.add(old, dx, out=dx)
            tape.put(x, dx)
    tape.record(bwd, (m.weight,))
    return y


def adaptive_avgpool(tape: Tape, x, out_hw):
    y = ops.adaptive_avgpool_fwd(x, out_hw)

    def bwd():
        dy = tape.pop(y)
        if dy is not None:
            tape.put(x, ops.adaptive_avgpool_bwd(dy, (x.shape[1], x.shape[2])))
    tape.record(bwd)
    return y


def bilinear_into(tape: Tape, x, out, align_corners):
    """Bilinear resize of x written into `out` (a channel slice of a concat buffer)."""
    ops.bilinear_fwd(x, (out.shape[1], out.shape[2]), align_corners, out=out)

    def bwd():
        dy = tape.pop(out)
        if dy is not None:
            tape.put(x, ops.bilinear_bwd(dy, (x.shape[1], x.shape[2]), align_corners))
    tape.record(bwd)
    return out


def concat(tape: Tape, buf, parts):
    """`buf` (N,H,W,sum C) already holds `parts` (its channel slices, in order, written in place by their producers):
    torch.cat without a copy. Backward hands every producer its slice of the buffer's gradient."""
    def bwd():
        d = tape.pop(buf)
        if d is None:
            return
        c0 = 0
        for part in parts:
            tape.put(part, d[..., c0:c0 + part.shape[3]])
            c0 += part.shape[3]
    tape.record(bwd)
    return buf


def logits_head(tape: Tape, x, n_classes, out_hw, align_corners):
    """NHWC bf16 (channel-padded) -> fp32 NCHW logits, bilinearly resized to out_hw when it differs
    (models/resunet.py:51-56; smp SegmentationHead upsampling). Returns (logits, feed_grad(g))."""
    h, w, cpad = x.shape[1], x.shape[2], x.shape[3]
    same = (h, w) == tuple(out_hw)
    if same:
        out = ops.nhwc_to_nchw(x, n_classes)
    else:
        out = ops.bilinear_head_fwd(x, n_classes, out_hw, align_corners)

    def feed(g):
        if g.dtype == torch.bfloat16:      # already the bf16 NHWC (N, h, w, cpad) gradient (ops.dice_bwd_nhwc)
            assert same and tuple(g.shape) == tuple(x.shape)
            dx = g
        elif same:
            dx = ops.nchw_to_nhwc(g.contiguous(), c_pad=cpad)
        else:
            dx = ops.bilinear_head_bwd(g, (h, w, cpad), align_corners)
        tape.put(x, dx)
    feed.nhwc_cpad = cpad if same else None     # callers may hand feed() a bf16 NHWC gradient with this channel padding
    return out, feed


def head_conv3x3(tape: Tape, m: Conv2d, x):
    """smp SegmentationHead at full resolution, Conv2d(C, classes, 3, padding=1) + bias -> fp32 NCHW logits, on the
    direct kernels of head.cu (fp32 master weights; no 16-channel padded tensors). -> (logits, feed_grad(g))."""
    w = m.weight.detach()
    bias = m.bias.detach() if m.bias is not None else None
    out = ops.head3x3_fwd(x, w, bias)

    def feed(g):
        g = g.contiguous()
        ops.head3x3_wgrad(x, g, tape.arena.slot(m.weight), tape.arena.slot(m.bias) if m.bias is not None else None)
        if tape.wants(x):
            old = tape.pop(x)
            dx = ops.head3x3_dgrad(g, w, tuple(x.shape))
            if old is not None:
                ops.add(old, dx, out=dx)
            tape.put(x, dx)
    feed.nhwc_cpad = None
    # feed() runs before the tape's backward steps: this (first executed) step only reports the head's parameters as
    # complete to the data-parallel reducer
    tape.record(lambda: None, (m.weight, m.bias))
    return out, feed


# ------------------------------------------------------------------------------------------------ ResNet18 encoder
def basic_block(tape: Tape, blk, x):
    """torchvision BasicBlock.forward (resnet.py:89-106): relu(bn2(conv2(relu(bn1(conv1 x)))) + shortcut(x))."""
    out = conv_bn_act(tape, blk.conv1, blk.bn1, [x], relu=True)
    if blk.downsample is not None:
        ds = blk.downsample
        identity = conv_bn_act(tape, getattr(ds, "0"), getattr(ds, "1"), [x], relu=False)
    else:
        identity = x
    return conv_bn_act(tape, blk.conv2, blk.bn2, [out], relu=True, residual=identity)


def stem_input(tape: Tape, x):
    """fp32 NCHW (B,3,H,W) -> the stem's space-to-depth bf16 input; a bf16 (B,H/2,W/2,16) tensor passes through
    (already prepared on the GPU from uint8 patches by ops.stem_prep_u8)."""
    if x.dtype == torch.bfloat16:
        z = x
    else:
        z = ops.stem_prep_f32(x)
    tape.no_grad_for(z)
    return z


def resnet18_features(tape: Tape, enc, x, depth=5):  # noqa: E302
    """-> [stem (64 @ H/2), layer1 (64 @ H/4), layer2 (128 @ H/8), layer3 (256 @ H/16), layer4 (512 @ H/32)][:depth]
    (models/resunet.py:25-33; smp ResNetEncoder.forward stages 1..depth)."""
    z = stem_input(tape, x)
    s = stem(tape, enc.conv1, enc.bn1, z)
    feats = [s]
    if depth >= 2:
        f = maxpool(tape, s)
        for li in range(1, depth):
            for blk in enc.blocks(li):
                f = basic_block(tape, blk, f)
            feats.append(f)
    return feats
