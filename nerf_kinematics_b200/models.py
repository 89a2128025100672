"""The two MLP layer graphs of the path as nn.Modules whose state_dict layouts equal the reference's.

* ``NeRF``              -- the Keras functional graph of init_nerf_model
                           (nerf/run_nerf_helpers.py:80-118; D x W trunk, skip after layer 4, alpha /
                           bottleneck / views / rgb heads).  Keys: pts_linears.{i}, alpha_linear,
                           bottleneck_linear, views_linears.0, rgb_linear | output_linear.
* ``FlexibleNeRFModel`` -- nerf-pytorch/nerf/models.py:185-256 at the depths that model can actually
                           run (num_layers <= 5, where its skip never fires).  Same constructor
                           signature, same parameter registration order (so optimizer state dicts of the
                           shipped checkpoints line up), same state_dict keys.

The fp32 master weights are ordinary nn.Parameters (any torch optimizer works) that alias one flat
device buffer; the kernels behind include/nerf_b200.h read that buffer (fp32 path) or a packed bf16
image of it that is refreshed whenever the parameters change (tcgen05 path).  The modules hold no
torch implementation of the network: evaluation goes through the CUDA library or raises.
"""
import ctypes as C
import math

import torch
import torch.nn as nn

from . import _lib
from ._lib import check, lib

GRAPH_NERF, GRAPH_FLEXIBLE = 0, 1
# "fp16" is what the tensor-core path computes in (IEEE half operands, fp32 accumulation); "bf16" is kept as an alias
# because north_star / the reference-facing kwargs call it the "bf16 tensor-core path".
_PREC = {"fp32": 0, "fp16": 1, "f16": 1, "bf16": 1, 0: 0, 1: 1}


def precision_id(p) -> int:
    try:
        return _PREC[p]
    except KeyError:
        raise ValueError("precision must be 'fp32' or 'fp16' (alias 'bf16'), got %r" % (p,))


class KernelMLP(nn.Module):
    """Common machinery: native handle, flat parameter buffer, bf16 repacking, point queries."""

    def __init__(self):
        super().__init__()
        self._handle = None
        self._flat = None
        self._flat_grad = None
        self._layout = None  # [(state_dict key, shape, offset)]
        self._packed_fp = {}
        self.precision = "fp32"

    # -- native handle -------------------------------------------------------------------------
    def _create_handle(self, kind, depth, width, n_freq_xyz, n_freq_dir, use_viewdirs, skip):
        h = C.c_void_p()
        check(lib().nb_model_create(kind, depth, width, n_freq_xyz, n_freq_dir, int(bool(use_viewdirs)), skip,
                                    C.byref(h)), "nb_model_create")
        self._handle = h
        n = lib().nb_model_num_tensors(h)
        layout = []
        own = dict(self.named_parameters())
        for i in range(n):
            name = C.create_string_buffer(128)
            rows, cols, off = C.c_int64(), C.c_int64(), C.c_int64()
            check(lib().nb_model_tensor_info(h, i, name, 128, C.byref(rows), C.byref(cols), C.byref(off)))
            key = name.value.decode()
            shape = (rows.value, cols.value) if cols.value > 0 else (rows.value,)
            if key not in own or tuple(own[key].shape) != shape:
                raise _lib.NerfB200Error("module/native layout mismatch for %s: %s vs %s"
                                         % (key, tuple(own[key].shape) if key in own else None, shape))
            layout.append((key, shape, off.value))
        if len(layout) != len(own):
            raise _lib.NerfB200Error("module has %d parameters, native graph %d" % (len(own), len(layout)))
        self._layout = layout
        self.n_params = int(lib().nb_model_param_count(h))
        self.flops_per_point = int(lib().nb_model_flops_per_point(h))

    def __del__(self):
        try:
            if self._handle is not None and _lib._lib is not None:
                _lib._lib.nb_model_destroy(self._handle)
        except Exception:  # interpreter shutdown
            pass

    # -- flat parameter buffer -----------------------------------------------------------------
    def _is_flat(self):
        f = self._flat
        if f is None:
            return False
        base = f.data_ptr()
        own = dict(self.named_parameters())
        for key, _, off in self._layout:
            p = own[key]
            if p.device != f.device or p.dtype != torch.float32 or p.data_ptr() != base + 4 * off:
                return False
        return True

    def flatten_parameters(self):
        """Alias every parameter to one flat fp32 device buffer in the native layout (idempotent)."""
        if self._is_flat():
            return self._flat
        own = dict(self.named_parameters())
        dev = next(iter(own.values())).device
        if dev.type != "cuda":
            raise _lib.NerfB200Error("model parameters are on %s: nerf_kinematics_b200 has no CPU fallback; "
                                     "call .cuda() first" % dev)
        flat = torch.empty((self.n_params,), dtype=torch.float32, device=dev)
        with torch.no_grad():
            for key, shape, off in self._layout:
                p = own[key]
                n = int(math.prod(shape))
                view = flat[off:off + n].view(shape)
                view.copy_(p.data)
                p.data = view
        self._flat = flat
        check(lib().nb_model_set_params(self._handle, C.c_void_p(flat.data_ptr()), self.n_params),
              "nb_model_set_params")
        self._packed_fp = {}
        return flat

    def flat_grad(self, zero=False):
        """A flat fp32 gradient buffer in the parameter layout (used by the train step / all-reduce)."""
        flat = self.flatten_parameters()
        if self._flat_grad is None or self._flat_grad.device != flat.device:
            self._flat_grad = torch.zeros_like(flat)
        elif zero:
            self._flat_grad.zero_()
        return self._flat_grad

    def grad_views(self, flat_grad):
        own = dict(self.named_parameters())
        out = []
        for key, shape, off in self._layout:
            n = int(math.prod(shape))
            out.append((own[key], flat_grad[off:off + n].view(shape)))
        return out

    def ordered_parameters(self):
        """Parameters in the native flat order (the order the render autograd function uses)."""
        own = dict(self.named_parameters())
        return [own[key] for key, _, _ in self._layout]

    def _fingerprint(self):
        return tuple(p._version for p in self.parameters()) + (self._flat.data_ptr(),)

    def mark_params_changed(self):
        """The flat buffer was written by a kernel (fused Adam): torch's version counters did not see it."""
        self._packed_fp = {}

    def prepare(self, precision=None, force=False):
        """Make the native handle current for `precision`; returns (handle, precision id).  `force` repacks the
        16-bit images unconditionally (the captured train step must not depend on host-side version checks)."""
        prec = precision_id(self.precision if precision is None else precision)
        self.flatten_parameters()
        if prec == 1:
            fp = self._fingerprint()
            if force or self._packed_fp.get(prec) != fp:
                with torch.cuda.device(self._flat.device):
                    check(lib().nb_model_pack(self._handle, prec,
                                              C.c_void_p(torch.cuda.current_stream().cuda_stream)), "nb_model_pack")
                self._packed_fp[prec] = fp
        return self._handle, prec

    # -- evaluation ----------------------------------------------------------------------------
    def query(self, pts, viewdirs=None, precision=None):
        """network_query_fn semantics (nerf/run_nerf.py:30-45): pts [..., S, 3], viewdirs [..., 3] -> raw [..., S, 4].
        A 2-D pts [N, 3] is N independent points: viewdirs must then be [N, 3] (one direction per point)."""
        from .ops import _f32, _optptr, _ptr, _stream
        h, prec = self.prepare(precision)
        p = _f32(pts, "pts")
        lead = p.shape[:-1]
        S = p.shape[-2] if p.dim() >= 3 else 1
        p2 = p.reshape(-1, 3)
        vd = None
        if self.use_viewdirs:
            if viewdirs is None:
                raise ValueError("this model was built with use_viewdirs=True")
            vd = _f32(viewdirs, "viewdirs").reshape(-1, 3)
            if vd.shape[0] * S != p2.shape[0]:
                raise ValueError("viewdirs [%d,3] do not match pts %s: pass one direction per ray for pts "
                                 "[N_rays, S, 3], one per point for pts [N, 3]" % (vd.shape[0], tuple(pts.shape)))
        n = p2.shape[0]
        raw = torch.empty((n, 4), dtype=torch.float32, device=p.device)
        with torch.cuda.device(p.device):
            nbytes = lib().nb_mlp_workspace_bytes(h, prec, n, 0)
            ws = torch.empty((max(nbytes, 1),), dtype=torch.uint8, device=p.device)
            check(lib().nb_mlp_forward_points(h, prec, _ptr(p2), _optptr(vd), n, S, _ptr(raw), _ptr(ws), nbytes,
                                              _stream()), "nb_mlp_forward_points")
        return raw.reshape(*lead, 4)

    def forward(self, x):
        raise NotImplementedError(
            "the B200 path fuses the positional encoding into the network kernel; call "
            "model.query(pts, viewdirs) or network_query_fn(pts, viewdirs, model) instead of model(embedded)")


class NeRF(KernelMLP):
    """init_nerf_model(D, W, input_ch, input_ch_views, skips=[4], use_viewdirs) -- nerf/run_nerf_helpers.py:80-118."""

    def __init__(self, D=8, W=256, input_ch=63, input_ch_views=27, output_ch=4, skips=(4,), use_viewdirs=True):
        super().__init__()
        if (input_ch - 3) % 6 or (use_viewdirs and (input_ch_views - 3) % 6):
            raise ValueError("input_ch / input_ch_views must be 3 + 6*L (include_input embedders)")
        skips = [s for s in skips if 0 <= s < D - 1]
        if len(skips) > 1:
            raise ValueError("one skip connection is supported (the reference uses skips=[4])")
        self.D, self.W, self.input_ch, self.input_ch_views = D, W, input_ch, input_ch_views
        self.skips, self.use_viewdirs = tuple(skips), bool(use_viewdirs)
        skip = skips[0] if skips else -1
        ins = [input_ch] + [(W + input_ch) if (i == skip) else W for i in range(D - 1)]
        self.pts_linears = nn.ModuleList([nn.Linear(ins[i], W) for i in range(D)])
        if use_viewdirs:
            self.alpha_linear = nn.Linear(W, 1)
            self.bottleneck_linear = nn.Linear(W, 256)
            self.views_linears = nn.ModuleList([nn.Linear(256 + input_ch_views, W // 2)])
            self.rgb_linear = nn.Linear(W // 2, 3)
        else:
            self.output_linear = nn.Linear(W, output_ch)
        # Keras Dense defaults: glorot_uniform kernel, zero bias
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                nn.init.zeros_(m.bias)
        self.n_freq_xyz, self.n_freq_dir = (input_ch - 3) // 6, ((input_ch_views - 3) // 6 if use_viewdirs else 0)
        self._create_handle(GRAPH_NERF, D, W, self.n_freq_xyz, self.n_freq_dir, use_viewdirs, skip)

    # Keras get_weights() order (np.save(net.get_weights()), nerf/run_nerf.py:835-843): [kernel[in,out], bias] per layer,
    # layers of a functional Model sorted by graph depth -- the trunk, then bottleneck, views, rgb, and alpha LAST
    # (alpha's Dense(1) feeds only the final concat, so it sits at the same depth as rgb and after it): index 2D =
    # bottleneck [W,256], 2D+2 = views [256+27, W/2], 2D+4 = rgb [W/2,3], 2D+6 = alpha [W,1] -- the layout of the
    # official model_200000.npy files and of the well-known load_weights_from_keras converters.
    def keras_layer_names(self):
        names = ["pts_linears.%d" % i for i in range(self.D)]
        names += ["bottleneck_linear", "views_linears.0", "rgb_linear", "alpha_linear"] if self.use_viewdirs \
            else ["output_linear"]
        return names

    def get_weights(self):
        sd = self.state_dict()
        out = []
        for n in self.keras_layer_names():
            out.append(sd[n + ".weight"].detach().t().contiguous().cpu().numpy())
            out.append(sd[n + ".bias"].detach().cpu().numpy())
        return out

    def set_weights(self, weights):
        names = self.keras_layer_names()
        if len(weights) != 2 * len(names):
            raise ValueError("expected %d arrays, got %d" % (2 * len(names), len(weights)))
        sd = {}
        own = self.state_dict()
        for i, n in enumerate(names):
            w, b = torch.as_tensor(weights[2 * i]), torch.as_tensor(weights[2 * i + 1])
            if tuple(w.shape) != tuple(own[n + ".weight"].shape[::-1]) or tuple(b.shape) != tuple(own[n + ".bias"].shape):
                raise ValueError("weights[%d] / [%d] have shapes %s / %s; %s expects kernel %s, bias %s (Keras order: "
                                 "trunk, bottleneck, views, rgb, alpha)" % (2 * i, 2 * i + 1, tuple(w.shape), tuple(b.shape),
                                                                            n, tuple(own[n + ".weight"].shape[::-1]),
                                                                            tuple(own[n + ".bias"].shape)))
            sd[n + ".weight"] = w.t().contiguous()
            sd[n + ".bias"] = b
        self.load_state_dict(sd)


class FlexibleNeRFModel(KernelMLP):
    """nerf-pytorch/nerf/models.py:185-256 (same signature, registration order and keys)."""

    def __init__(self, num_layers=4, hidden_size=128, skip_connect_every=4, num_encoding_fn_xyz=6,
                 num_encoding_fn_dir=4, include_input_xyz=True, include_input_dir=True, use_viewdirs=True):
        super().__init__()
        if not include_input_xyz or (use_viewdirs and not include_input_dir):
            raise NotImplementedError("include_input=False embedders are not used by any reference config")
        if num_layers - 2 >= skip_connect_every > 0 and num_layers - 1 > skip_connect_every:
            # the reference forward reads self.linear_layers (models.py:243) and raises AttributeError here
            raise NotImplementedError("FlexibleNeRFModel with an active skip connection crashes in the reference "
                                      "(models.py:243); use NeRF for the 8x256 graph")
        self.dim_xyz = 3 + 6 * num_encoding_fn_xyz
        self.dim_dir = (3 + 6 * num_encoding_fn_dir) if use_viewdirs else 0
        self.skip_connect_every = skip_connect_every
        self.layer1 = nn.Linear(self.dim_xyz, hidden_size)
        self.layers_xyz = nn.ModuleList([nn.Linear(hidden_size, hidden_size) for _ in range(num_layers - 1)])
        self.use_viewdirs = bool(use_viewdirs)
        if use_viewdirs:
            self.layers_dir = nn.ModuleList([nn.Linear(self.dim_dir + hidden_size, hidden_size // 2)])
            self.fc_alpha = nn.Linear(hidden_size, 1)
            self.fc_rgb = nn.Linear(hidden_size // 2, 3)
            self.fc_feat = nn.Linear(hidden_size, hidden_size)
        else:
            self.fc_out = nn.Linear(hidden_size, 4)
        self.n_freq_xyz, self.n_freq_dir = num_encoding_fn_xyz, (num_encoding_fn_dir if use_viewdirs else 0)
        self._create_handle(GRAPH_FLEXIBLE, num_layers, hidden_size, self.n_freq_xyz, self.n_freq_dir, use_viewdirs, -1)


def model_from_state_dict(sd, use_viewdirs=None):
    """Build the right module for a reference state_dict (either key family) and load it."""
    keys = set(sd.keys())
    if "layer1.weight" in keys:
        dim_xyz = sd["layer1.weight"].shape[1]
        hidden = sd["layer1.weight"].shape[0]
        n_xyz = len([k for k in keys if k.startswith("layers_xyz.") and k.endswith(".weight")])
        uv = "fc_alpha.weight" in keys
        n_dir = (sd["layers_dir.0.weight"].shape[1] - hidden - 3) // 6 if uv else 0
        m = FlexibleNeRFModel(num_layers=n_xyz + 1, hidden_size=hidden, num_encoding_fn_xyz=(dim_xyz - 3) // 6,
                              num_encoding_fn_dir=n_dir, use_viewdirs=uv)
    elif "pts_linears.0.weight" in keys:
        D = len([k for k in keys if k.startswith("pts_linears.") and k.endswith(".weight")])
        W, input_ch = sd["pts_linears.0.weight"].shape
        uv = "alpha_linear.weight" in keys
        skips = [i for i in range(D - 1) if sd["pts_linears.%d.weight" % (i + 1)].shape[1] == W + input_ch]
        views = sd["views_linears.0.weight"].shape[1] - 256 if uv else 3
        m = NeRF(D=D, W=W, input_ch=input_ch, input_ch_views=views, skips=skips, use_viewdirs=uv)
    else:
        raise ValueError("unrecognised state_dict keys: %s" % sorted(keys)[:4])
    m.load_state_dict({k: torch.as_tensor(v) for k, v in sd.items()})
    return m
