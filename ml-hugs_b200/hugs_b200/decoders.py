"""The three decoder MLPs of the human model on the tensor cores.

Mirrors hugs/models/modules/decoders.py:24-110 -- same class names, constructor arguments, sub-module
structure and therefore the same ``state_dict`` keys (a reference checkpoint loads) and the same result
dictionaries -- with ``forward`` running on ``hugs_b200_mlp_chains_forward`` (csrc/mlp_tc.cu: one CTA keeps a
128-row tile on the SM through the whole decoder, tcgen05.mma kind::tf32 with the 3xTF32 split, fp32
accumulators in TMEM, exact GELU in the epilogue).  ``decode_all`` evaluates the three decoders of
hugs/models/hugs_trimlp.py:205-238 on the tri-plane features in ONE launch.

Backward: the forward kernel saves every layer's pre-activations and ``hugs_b200_mlp_chains_backward`` runs
the whole chain backwards on the same tiles: dA = dZ W against the packed W^T, dW^T = A^T dZ from transposed
K-major tiles, db by warp shuffles, GELU' in the epilogue.
Only ``act='gelu'`` (the reference's default and the release configuration) takes the fused path; the other
activations (and the pose-corrective blend shapes, off in the release configuration) run the module's own
arithmetic on the GPU.  There is no CPU path: CPU inputs raise.
"""
import ctypes as C
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _native as N

_ACTS = {'softplus': nn.Softplus, 'relu': nn.ReLU, 'gelu': nn.GELU, 'tanh': nn.Tanh}


def _act_module(act):
    if act not in _ACTS:
        raise ValueError("activation %r is not supported (the reference's 'sine' needs its SineActivation)" % (act,))
    return _ACTS[act]()


def _require_cuda(x):
    if not x.is_cuda:
        raise RuntimeError("hugs_b200 has no CPU path: decoder inputs must be CUDA tensors")


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _pad(n, hidden=True):
    # every layer's width is rounded up to 32: the backward kernel walks the widths in 32-element K blocks
    return (n + 31) // 32 * 32


class _Spec:
    """One chain: hidden layers [(W, b)], the head (W rows concatenated over the decoder's heads, b) and the
    head's output slices [(name, col0, ncols, act)]."""

    def __init__(self, hidden, head, slices):
        self.hidden, self.head, self.slices = hidden, head, slices



class _MlpChains(torch.autograd.Function):
    """forward(ctx, x, specs_meta, *tensors): tensors = per chain, per layer (W, b) in order (head last)."""

    @staticmethod
    def forward(ctx, x, meta, *params):
        if not x.is_cuda:
            raise RuntimeError("hugs_b200 has no CPU path: decoder inputs must be CUDA tensors")
        device = x.device
        x = x.float().contiguous()
        rows, Kx = int(x.shape[0]), int(x.shape[1])
        need_grad = any(ctx.needs_input_grad)   # all False under torch.no_grad(): no pre-activations are kept
        chains = (N.MlpChain * len(meta))()
        keep, outs, zs = [], [], []
        p = 0
        with torch.cuda.device(device):
            for c, (n_layers, slices) in enumerate(meta):
                ch = chains[c]
                ch.n_layers, ch.n_outs = n_layers, len(slices)
                K = Kx
                zc = []
                for l in range(n_layers):
                    W, b = params[p].detach().float().contiguous(), params[p + 1].detach().float().contiguous()
                    p += 2
                    n_true = int(W.shape[0])
                    n_pad = _pad(n_true)
                    packed = torch.empty(int(N.mlp_packed_floats(K, n_pad)), dtype=torch.float32, device=device)
                    N.check(N.mlp_pack(W.data_ptr(), b.data_ptr(), n_true, n_pad, K, packed.data_ptr(), _stream(device)),
                            "mlp_pack")
                    z = torch.empty(rows, n_pad, dtype=torch.float32, device=device) if need_grad else None
                    keep += [W, b, packed]
                    zc.append(z)
                    ch.layer[l].K, ch.layer[l].N = K, n_pad
                    ch.layer[l].act = N.MLP_ACT_GELU
                    ch.layer[l].packed = packed.data_ptr()
                    ch.layer[l].z_out = z.data_ptr() if z is not None else None
                    K = n_pad
                oc = []
                for s, (_, col0, ncols, act) in enumerate(slices):
                    o = torch.empty(rows, ncols, dtype=torch.float32, device=device)
                    ch.out[s].ptr, ch.out[s].ld, ch.out[s].col0, ch.out[s].ncols, ch.out[s].act = (
                        o.data_ptr(), ncols, col0, ncols, act)
                    oc.append(o)
                outs.append(oc)
                zs.append(zc)
            if rows > 0:
                N.check(N.mlp_chains_forward(rows, x.data_ptr(), Kx, Kx, len(meta), chains, _stream(device)),
                        "mlp_chains_forward")
        flat_out = [o for oc in outs for o in oc]
        if need_grad:
            ctx.meta = meta
            ctx.n_params = len(params)
            ctx.save_for_backward(x, *[t.detach() for t in params], *[z for zc in zs for z in zc], *flat_out)
        return tuple(flat_out)

    @staticmethod
    def backward(ctx, *grads):
        meta, n_params = ctx.meta, ctx.n_params
        saved = ctx.saved_tensors
        x, params = saved[0], saved[1:1 + n_params]
        n_z = sum(n for n, _ in meta)
        zs = saved[1 + n_params:1 + n_params + n_z]
        device = x.device
        rows, Kx = int(x.shape[0]), int(x.shape[1])
        chains = (N.MlpBwdChain * len(meta))()
        keep, dWts, dbs, dxs = [], [], [], []
        p = zi = oi = 0
        with torch.cuda.device(device):
            for c, (n_layers, slices) in enumerate(meta):
                ch = chains[c]
                ch.n_layers, ch.n_outs = n_layers, len(slices)
                K = Kx
                for l in range(n_layers):
                    W = params[p + 2 * l].float().contiguous()
                    n_true, n_pad = int(W.shape[0]), _pad(int(W.shape[0]))
                    packed_t = torch.empty(int(N.mlp_packed_floats(n_pad, K)), dtype=torch.float32, device=device)
                    N.check(N.mlp_pack_t(W.data_ptr(), n_true, K, n_pad, packed_t.data_ptr(), _stream(device)), "mlp_pack_t")
                    dWt = torch.zeros(K, n_pad, dtype=torch.float32, device=device)
                    db = torch.zeros(n_pad, dtype=torch.float32, device=device)
                    keep += [W, packed_t]
                    dWts.append((dWt, n_true))
                    dbs.append((db, n_true))
                    L = ch.layer[l]
                    L.K, L.N, L.packed_t = K, n_pad, packed_t.data_ptr()
                    L.z_prev = zs[zi + l - 1].data_ptr() if l > 0 else None
                    L.dWt, L.db = dWt.data_ptr(), db.data_ptr()
                    K = n_pad
                ch.z_head = zs[zi + n_layers - 1].data_ptr()
                dx = torch.empty(rows, Kx, dtype=torch.float32, device=device)
                dxs.append(dx)
                ch.dx = dx.data_ptr()
                for s, (_, col0, ncols, act) in enumerate(slices):
                    g = grads[oi + s]
                    if g is not None:
                        g = g.float().contiguous()
                        keep.append(g)
                    ch.out[s].dptr = g.data_ptr() if g is not None else None
                    ch.out[s].ld, ch.out[s].col0, ch.out[s].ncols, ch.out[s].act = ncols, col0, ncols, act
                p += 2 * n_layers
                zi += n_layers
                oi += len(slices)
            if rows > 0:
                N.check(N.mlp_chains_backward(rows, x.data_ptr(), Kx, Kx, len(meta), chains, _stream(device)),
                        "mlp_chains_backward")
        dparams = []
        for (dWt, n_true), (db, _) in zip(dWts, dbs):
            dparams += [dWt.t()[:n_true].contiguous(), db[:n_true]]
        dx = dxs[0] if len(dxs) == 1 else torch.stack(dxs).sum(0)
        if rows == 0:
            dx = torch.zeros_like(x)
        return (dx, None, *dparams)


def run_chains(x, specs):
    """specs: list of _Spec.  Returns, per chain, {slice name: tensor}."""
    meta, params = [], []
    for sp in specs:
        layers = list(sp.hidden) + [sp.head]
        meta.append((len(layers), tuple(sp.slices)))
        for W, b in layers:
            params += [W, b]
    flat = _MlpChains.apply(x, tuple(meta), *params)
    out, i = [], 0
    for sp in specs:
        out.append({name: flat[i + s] for s, (name, _, _, _) in enumerate(sp.slices)})
        i += len(sp.slices)
    return out


def _lin(m):
    return (m.weight, m.bias)


class AppearanceDecoder(nn.Module):
    """hugs/models/modules/decoders.py:24-44."""

    def __init__(self, n_features, hidden_dim=64, act='gelu'):
        super().__init__()
        self.hidden_dim, self.act = hidden_dim, act
        self.net = nn.Sequential(nn.Linear(n_features, hidden_dim), _act_module(act),
                                 nn.Linear(hidden_dim, hidden_dim), _act_module(act))
        self.opacity = nn.Sequential(nn.Linear(hidden_dim, 1), nn.Sigmoid())
        self.shs = nn.Linear(hidden_dim, 16 * 3)

    def spec(self):
        # head columns: shs first so that its 192-byte rows stay 16-byte aligned
        W = torch.cat([self.shs.weight, self.opacity[0].weight], 0)
        b = torch.cat([self.shs.bias, self.opacity[0].bias], 0)
        return _Spec([_lin(self.net[0]), _lin(self.net[2])], (W, b),
                     [("shs", 0, 48, N.MLP_ACT_NONE), ("opacity", 48, 1, N.MLP_ACT_SIGMOID)])

    def forward(self, x):
        _require_cuda(x)
        if self.act != 'gelu':   # not the release configuration: the module's own arithmetic, on the GPU
            h = self.net(x)
            return {'shs': self.shs(h), 'opacity': self.opacity(h)}
        return run_chains(x, [self.spec()])[0]


class DeformationDecoder(nn.Module):
    """hugs/models/modules/decoders.py:47-84."""

    def __init__(self, n_features, hidden_dim=128, weight_norm=True, act='gelu', disable_posedirs=False):
        super().__init__()
        self.hidden_dim, self.act, self.disable_posedirs = hidden_dim, act, disable_posedirs
        self.net = nn.Sequential(nn.Linear(n_features, hidden_dim), _act_module(act),
                                 nn.Linear(hidden_dim, hidden_dim), _act_module(act))
        self.skinning_linear = nn.Linear(hidden_dim, hidden_dim)
        self.skinning = nn.Linear(hidden_dim, 24)
        if weight_norm:
            self.skinning_linear = nn.utils.weight_norm(self.skinning_linear)
        if not disable_posedirs:
            self.blendshapes = nn.Linear(hidden_dim, 3 * 207)
            nn.init.constant_(self.blendshapes.bias, 0.0)
            nn.init.constant_(self.blendshapes.weight, 0.0)

    def _skinning_linear_weight(self):
        m = self.skinning_linear
        if hasattr(m, "weight_g"):   # torch.nn.utils.weight_norm: w = g * v / ||v|| per output row
            return torch._weight_norm(m.weight_v, m.weight_g, 0)
        return m.weight

    def spec(self):
        return _Spec([_lin(self.net[0]), _lin(self.net[2]), (self._skinning_linear_weight(), self.skinning_linear.bias)],
                     _lin(self.skinning), [("lbs_weights", 0, 24, N.MLP_ACT_GELU)])

    def forward(self, x):
        _require_cuda(x)
        if self.act != 'gelu' or not self.disable_posedirs:
            h = self.net(x)
            posedirs = None
            if not self.disable_posedirs:
                posedirs = self.blendshapes(h).reshape(207, -1)
            lbs_weights = F.gelu(self.skinning(F.gelu(self.skinning_linear(h))))
            return {'lbs_weights': lbs_weights, 'posedirs': posedirs}
        return {'lbs_weights': run_chains(x, [self.spec()])[0]['lbs_weights'], 'posedirs': None}


class GeometryDecoder(nn.Module):
    """hugs/models/modules/decoders.py:87-110 (the three heads share ``net``)."""

    def __init__(self, n_features, use_surface=False, hidden_dim=128, act='gelu'):
        super().__init__()
        self.hidden_dim, self.act = hidden_dim, act
        self.net = nn.Sequential(nn.Linear(n_features, hidden_dim), _act_module(act),
                                 nn.Linear(hidden_dim, hidden_dim), _act_module(act))
        self.xyz = nn.Sequential(self.net, nn.Linear(hidden_dim, 3))
        self.rotations = nn.Sequential(self.net, nn.Linear(hidden_dim, 6))
        self.scales = nn.Sequential(self.net, nn.Linear(hidden_dim, 2 if use_surface else 3))

    def spec(self):
        heads = [self.xyz[1], self.rotations[1], self.scales[1]]
        W = torch.cat([h.weight for h in heads], 0)
        b = torch.cat([h.bias for h in heads], 0)
        ns = int(self.scales[1].weight.shape[0])
        return _Spec([_lin(self.net[0]), _lin(self.net[2])], (W, b),
                     [("xyz", 0, 3, N.MLP_ACT_NONE), ("rotations", 3, 6, N.MLP_ACT_NONE),
                      ("scales", 9, ns, N.MLP_ACT_GELU)])

    def forward(self, x):
        _require_cuda(x)
        if self.act != 'gelu':
            return {'xyz': self.xyz(x), 'rotations': self.rotations(x), 'scales': F.gelu(self.scales(x))}
        return run_chains(x, [self.spec()])[0]


def decode_all(tri_feats, appearance_dec, geometry_dec, deformation_dec=None):
    """The decoder part of hugs/models/hugs_trimlp.py:205-238 (canon_forward) in ONE launch:
    returns (appearance_out, geometry_out, deformation_out | None) with the reference's keys."""
    decs = [appearance_dec, geometry_dec] + ([deformation_dec] if deformation_dec is not None else [])
    _require_cuda(tri_feats)
    fused = all(d.act == 'gelu' for d in decs) and \
        (deformation_dec is None or deformation_dec.disable_posedirs)
    if not fused:
        return (appearance_dec(tri_feats), geometry_dec(tri_feats),
                deformation_dec(tri_feats) if deformation_dec is not None else None)
    outs = run_chains(tri_feats, [d.spec() for d in decs])
    deform = None
    if deformation_dec is not None:
        deform = {'lbs_weights': outs[2]['lbs_weights'], 'posedirs': None}
    return outs[0], outs[1], deform


def canon_forward(triplane, xyz, appearance_dec, geometry_dec, deformation_dec=None, scaling_multiplier=1.0):
    """hugs/models/hugs_trimlp.py:205-238 (``canon_forward``) on this repository's kernels: tri-plane features
    (``hugs_b200.triplane.TriPlane``: one gather kernel) -> the three decoders (ONE tcgen05 launch) -> the
    reference's result dictionary (``xyz_offsets``, ``scales``, ``rot6d_canon``, ``shs`` [N,16,3], ``opacity``,
    ``lbs_weights`` after softmax(./0.1), ``posedirs``).  ``hugs_b200.deform_gaussians`` takes it from there
    (hugs_trimlp.py:470-518)."""
    tri_feats = triplane(xyz)
    appearance_out, geometry_out, deformation_out = decode_all(tri_feats, appearance_dec, geometry_dec, deformation_dec)
    lbs_weights = posedirs = None
    if deformation_out is not None:
        lbs_weights = F.softmax(deformation_out['lbs_weights'] / 0.1, dim=-1)
        posedirs = deformation_out['posedirs']
    return {
        'xyz_offsets': geometry_out['xyz'],
        'scales': geometry_out['scales'] * scaling_multiplier,
        'rot6d_canon': geometry_out['rotations'],
        'shs': appearance_out['shs'].reshape(-1, 16, 3),
        'opacity': appearance_out['opacity'],
        'lbs_weights': lbs_weights,
        'posedirs': posedirs,
    }
