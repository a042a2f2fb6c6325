"""Host-side glue between torch tensors and the C ABI (device memory, streams: plumbing only)."""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import lib, check


def _np32(t):
    return np.ascontiguousarray(t.detach().to("cpu", torch.float32).numpy())


def _host_arrays(sd):
    """name -> contiguous fp32 numpy array.  CUDA tensors that are views of one storage (the Trainer's flat parameter
    buffer) come over in ONE device-to-host copy instead of one synchronising copy per tensor."""
    out, cache = {}, {}
    for k, v in sd.items():
        t = v.detach()
        if t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() and t.numel() > 0:
            st = t.untyped_storage()
            key = (t.device.index, st.data_ptr())
            if key not in cache:
                whole = torch.empty(0, dtype=torch.float32, device=t.device).set_(st, 0, (st.nbytes() // 4,))
                cache[key] = whole.cpu().numpy()
            off = t.storage_offset()
            out[k] = np.ascontiguousarray(cache[key][off:off + t.numel()].reshape(tuple(t.shape)))
        else:
            out[k] = _np32(t)
    return out


class _DevArray:
    """Stands in for a numpy array in Engine._pack: .shape and a device pointer."""

    def __init__(self, t):
        self.t, self.shape = t, tuple(t.shape)


def _ptr(a):
    if isinstance(a, _DevArray):
        return C.c_void_p(a.t.data_ptr())
    return a.ctypes.data_as(C.c_void_p)


def stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError("stofnet_b200: %s must be a CUDA tensor - this build has no CPU path (got %s)"
                           % (what, t.device))


class Engine:
    """Owns one stof_model handle and a grow-only workspace for one nn.Module instance."""

    def __init__(self, kind):
        self.kind = kind
        self.handle = None
        self.device = None
        self.key = None
        self.ws = None
        self.chunk_rows = 0     # 0 = library default
        self.path = None        # None = library default (tensor-core path for StofNet)
        self.layers_per_launch = 0

    def close(self):
        if self.handle is not None:
            lib.stof_model_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- weights ---------------------------------------------------------------------------
    def _pack(self, sd, on_device=False):
        """sd: mapping with the reference's state_dict key names -> ctypes struct (+ keep-alive).  on_device: the struct
        carries the DEVICE pointers of the (contiguous fp32 CUDA) tensors instead of host copies."""
        W = _lib.Weights()
        if on_device:
            keep = {}
            for k, v in sd.items():
                t = v.detach()
                if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
                    raise RuntimeError("stofnet_b200: device-side weight sync needs contiguous fp32 CUDA parameters (%s)" % k)
                keep[k] = _DevArray(t)
        else:
            keep = _host_arrays(sd)
        W.kind = self.kind
        if self.kind == _lib.KIND_EDSR:
            # models/edsr_1d.py:22-30
            nb = 1 + max(int(k.split(".")[1]) for k in keep if k.startswith("residual_blocks."))
            if 2 * nb + 1 > _lib.MAX_MID:
                raise ValueError("stofnet_b200: EDSR_1D with %d residual blocks (at most %d)" % (nb, (_lib.MAX_MID - 1) // 2))
            W.conv1_w, W.conv1_b = _ptr(keep["conv_input.weight"]), _ptr(keep["conv_input.bias"])
            for i in range(nb):
                for j, c in enumerate(("conv1", "conv2")):
                    W.mid_w[2 * i + j] = _ptr(keep["residual_blocks.%d.%s.weight" % (i, c)])
                    W.mid_b[2 * i + j] = _ptr(keep["residual_blocks.%d.%s.bias" % (i, c)])
            W.mid_w[2 * nb], W.mid_b[2 * nb] = _ptr(keep["conv_mid.weight"]), _ptr(keep["conv_mid.bias"])
            W.n_mid = 2 * nb + 1
            W.last_w, W.last_b = _ptr(keep["conv_output.weight"]), _ptr(keep["conv_output.bias"])
            W.num_features = keep["conv_input.weight"].shape[0]
            W.upsample_factor = W.num_features // keep["conv_output.weight"].shape[1]
            W.k_first = keep["conv_input.weight"].shape[2]
            W.k_mid = keep["conv_mid.weight"].shape[2]
            W.k_last = keep["conv_output.weight"].shape[2]
            W.sgb_scale = 1
            return W, keep
        W.conv1_w, W.conv1_b = _ptr(keep["conv1.weight"]), _ptr(keep["conv1.bias"])
        if self.kind == _lib.KIND_ESPCN:
            W.n_mid = 1
            W.mid_w[0], W.mid_b[0] = _ptr(keep["conv2.weight"]), _ptr(keep["conv2.bias"])
            W.last_w, W.last_b = _ptr(keep["conv3.weight"]), _ptr(keep["conv3.bias"])
            W.upsample_factor = keep["conv3.weight"].shape[0]
            W.num_features = 64
            W.k_first, W.k_mid, W.k_last = 5, 3, 3
            W.sgb_scale = 1
            return W, keep
        mids = sorted(int(k[4:-7]) for k in keep
                      if k.startswith("conv") and k.endswith(".weight") and k[4:-7].isdigit() and k != "conv1.weight")
        W.n_mid = len(mids)
        for j, i in enumerate(mids):
            W.mid_w[j], W.mid_b[j] = _ptr(keep["conv%d.weight" % i]), _ptr(keep["conv%d.bias" % i])
        W.last_w, W.last_b = _ptr(keep["conv_last.weight"]), _ptr(keep["conv_last.bias"])
        W.num_features = keep["conv1.weight"].shape[0]
        W.upsample_factor = keep["conv_last.weight"].shape[0]
        W.k_first = keep["conv1.weight"].shape[2]
        W.k_mid = keep["conv2.weight"].shape[2]
        W.k_last = keep["conv_last.weight"].shape[2]
        p = "semi_global_block."
        if p + "contract_conv.weight" in keep:
            cw = keep[p + "contract_conv.weight"]
            W.contract_w, W.contract_b = _ptr(cw), _ptr(keep[p + "contract_conv.bias"])
            W.expand_w, W.expand_b = _ptr(keep[p + "expand_conv.weight"]), _ptr(keep[p + "expand_conv.bias"])
            W.sgb_feat, W.sgb_k = cw.shape[0], cw.shape[2]
            W.sgb_scale = sd.sgb_scale if hasattr(sd, "sgb_scale") else 10 * (cw.shape[0] // cw.shape[1])
        else:
            W.sgb_scale, W.sgb_feat, W.sgb_k = 1, 0, 0
        return W, keep

    def sync(self, module, device):
        """(Re)upload the module's parameters when they changed (load_state_dict, optimizer step)."""
        params = dict(module.state_dict(keep_vars=True))
        key = (device.index, tuple((k, v.data_ptr(), v._version) for k, v in params.items()))
        if self.handle is not None and key == self.key:
            return
        W, keep = self._pack(params)
        if getattr(module, "semi_global_block", None) is not None:
            W.sgb_scale = int(module.semi_global_block.sample_scale)
        if self.handle is not None and self.device == device:
            check(lib.stof_model_set_weights(self.handle, C.byref(W)))
        else:
            self.close()
            h = C.c_void_p()
            check(lib.stof_model_create(C.byref(W), device.index if device.index is not None else 0, C.byref(h)))
            self.handle, self.device = h, device
            self._applied_path = None
        self.key = key
        del keep

    def sync_device(self, module, device, cache_key=None):
        """After an in-place optimizer step on the device (Trainer): re-derive the kernels' weight layouts from the
        parameters where they live, with one kernel on the current stream (stof_model_set_weights_device).
        cache_key (Trainer: the address of its flat parameter buffer, which every parameter is a view of): the pointer table
        is built once and reused while the key is the same - the walk over the state_dict was ~10 % of a batch-4 step."""
        cached = getattr(self, "_dev_pack", None)
        if cache_key is not None and cached is not None and cached[0] == (cache_key, id(module)):
            W = cached[1]
        else:
            params = dict(module.state_dict(keep_vars=True))
            W, keep = self._pack(params, on_device=True)
            if getattr(module, "semi_global_block", None) is not None:
                W.sgb_scale = int(module.semi_global_block.sample_scale)
            self._dev_pack = ((cache_key, id(module)), W, keep) if cache_key is not None else None
        with torch.cuda.device(device):
            check(lib.stof_model_set_weights_device(self.handle, C.byref(W), stream_ptr(device)))

    def set_path(self, path, layers_per_launch=0):
        """path: _lib.PATH_TC (tcgen05, fp16 hi/lo split operands) or _lib.PATH_SIMT (fp32 FFMA)."""
        self.path, self.layers_per_launch = path, layers_per_launch
        self._applied_path = None

    def _apply_path(self):
        want = (self.path, self.layers_per_launch)
        if self.path is not None and getattr(self, "_applied_path", None) != want:
            check(lib.stof_model_set_path(self.handle, int(self.path), int(self.layers_per_launch)))
            self._applied_path = want

    def chain_segment(self, module, first, count, in_planes, x0_planes=None):
        """Test hook: layers [first, first+count) of conv2..conv{nb-1}, conv_last on split fp16 planes
        (N, 2, 8, L, 8); returns planes, or the (N, 1, r*L) mask when the range ends with conv_last."""
        require_cuda(in_planes, "in_planes")
        dev = in_planes.device
        with torch.cuda.device(dev):
            self.sync(module, dev)
        N, _, _, L, _ = in_planes.shape
        n_tc = module.num_blocks - 1
        ends = first + count == n_tc
        in_planes = in_planes.contiguous()
        if ends:
            out = torch.empty((N, 1, module.upsample_factor * L), dtype=torch.float32, device=dev)
        else:
            out = torch.empty_like(in_planes)
        with torch.cuda.device(dev):
            check(lib.stof_chain_segment(self.handle, first, count, in_planes.data_ptr(),
                                         x0_planes.data_ptr() if x0_planes is not None else None, out.data_ptr(),
                                         N, L, stream_ptr(dev)))
        return out

    # -- forward ---------------------------------------------------------------------------
    def forward(self, module, x):
        require_cuda(x, "input")
        if x.dim() != 3 or x.shape[1] != 1:
            raise RuntimeError("stofnet_b200: expected input of shape (N, 1, L), got %s" % (tuple(x.shape),))
        dev = x.device
        x = x.detach().to(torch.float32).contiguous()
        N, _, L = x.shape
        r = module.upsample_factor
        with torch.cuda.device(dev):
            self.sync(module, dev)
            self._apply_path()
            need = C.c_size_t()
            check(lib.stof_forward_workspace_bytes(self.handle, N, L, self.chunk_rows, C.byref(need)))
            if self.ws is None or self.ws.numel() < need.value or self.ws.device != dev:
                self.ws = None
                self.ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
            out = torch.empty((N, 1, r * L), dtype=torch.float32, device=dev)
            check(lib.stof_forward(self.handle, x.data_ptr(), N, L, out.data_ptr(), self.ws.data_ptr(),
                                   self.ws.numel(), self.chunk_rows, stream_ptr(dev)))
        return out

    # -- training step -------------------------------------------------------------------
    def forward_train(self, module, x, synced=False):
        """Forward that keeps the activations in a workspace for backward(); returns (mask, x_contig, workspace).
        synced: the caller vouches that the handle holds the module's current parameters (Trainer, which tracks their versions)."""
        require_cuda(x, "input")
        if x.dim() != 3 or x.shape[1] != 1:
            raise RuntimeError("stofnet_b200: expected input of shape (N, 1, L), got %s" % (tuple(x.shape),))
        dev = x.device
        x = x.detach().to(torch.float32).contiguous()
        N, _, L = x.shape
        with torch.cuda.device(dev):
            if not (synced and self.handle is not None):
                self.sync(module, dev)
            self._apply_path()
            need = C.c_size_t()
            check(lib.stof_train_workspace_bytes(self.handle, N, L, C.byref(need)))
            ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
            out = torch.empty((N, 1, module.upsample_factor * L), dtype=torch.float32, device=dev)
            check(lib.stof_forward_train(self.handle, x.data_ptr(), N, L, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                         stream_ptr(dev)))
        return out, x, ws

    def backward(self, module, x, g_mask, ws, names, reuse=False):
        """d loss / d masks_pred -> one flat fp32 gradient buffer; returns (flat, [view per name in `names`]).
        reuse (Trainer only): the same gradient bucket, views and pointer table on every call - the kernels overwrite every
        gradient, so nothing needs clearing; the autograd path must NOT reuse (torch may keep the returned views as .grad)."""
        dev = x.device
        N, _, L = x.shape
        g_mask = g_mask.detach().to(torch.float32).contiguous()
        cached = getattr(self, "_grad_bucket", None) if reuse else None
        if cached is not None and cached[0] == (id(module), tuple(names), dev):
            _, flat, views, G = cached
        else:
            params = dict(module.named_parameters())
            sizes = [params[k].numel() for k in names]
            pad = [(-s) % 4 for s in sizes]                     # keep every view 16-byte aligned
            flat = torch.zeros((sum(sizes) + sum(pad),), dtype=torch.float32, device=dev)
            views, G, o = [], _lib.Grads(), 0
            for k, s, p in zip(names, sizes, pad):
                v = flat[o:o + s].view(params[k].shape)
                views.append(v)
                o += s + p
                self._set_grad_ptr(G, k, v.data_ptr())
            if reuse:
                self._grad_bucket = ((id(module), tuple(names), dev), flat, views, G)
        with torch.cuda.device(dev):
            check(lib.stof_backward(self.handle, x.data_ptr(), g_mask.data_ptr(), N, L, ws.data_ptr(), ws.numel(),
                                    C.byref(G), stream_ptr(dev)))
        return flat, views

    @staticmethod
    def _set_grad_ptr(G, key, ptr):
        stem, kind = key.rsplit(".", 1)
        sfx = "_w" if kind == "weight" else "_b"
        if stem == "conv1":
            setattr(G, "conv1" + sfx, ptr)
        elif stem == "conv_last":
            setattr(G, "last" + sfx, ptr)
        elif stem == "semi_global_block.contract_conv":
            setattr(G, "contract" + sfx, ptr)
        elif stem == "semi_global_block.expand_conv":
            setattr(G, "expand" + sfx, ptr)
        elif stem.startswith("conv") and stem[4:].isdigit():
            getattr(G, "mid" + sfx)[int(stem[4:]) - 2] = ptr
        else:
            raise KeyError("stofnet_b200: no gradient kernel for parameter %r" % key)

    def infer_host(self, module, x_host, window_size, threshold, cap=64, want_mask=False, device=None):
        """End-to-end eval step on HOST tensors (pinned recommended): H2D + forward + peaks + D2H."""
        dev = device if device is not None else next(module.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("stofnet_b200: module must live on a CUDA device (no CPU path)")
        x_host = x_host.detach().to(torch.float32).contiguous()
        if x_host.dim() == 3:
            x_host = x_host[:, 0, :]
        N, L = x_host.shape
        r = module.upsample_factor
        none = (threshold is None) or (not threshold)
        with torch.cuda.device(dev):
            self.sync(module, dev)
            self._apply_path()
            while True:
                # pinned straight from torch's caching host allocator: after the first call these are recycled blocks, not a
                # cudaHostAlloc + copy per call (which was ~40 % of the single-waveform latency)
                coords = torch.empty((N, cap), dtype=torch.float32, pin_memory=True)
                counts = torch.empty((N,), dtype=torch.int32, pin_memory=True)
                mask = torch.empty((N, 1, r * L), dtype=torch.float32, pin_memory=True) if want_mask else None
                kmax = C.c_int(0)
                check(lib.stof_infer_host(self.handle, x_host.data_ptr(), N, L, int(window_size), int(none),
                                          0.0 if none else float(threshold), coords.data_ptr(), cap, counts.data_ptr(),
                                          mask.data_ptr() if want_mask else None, C.byref(kmax)))
                if kmax.value <= cap:
                    break
                cap = kmax.value
        return coords, counts, kmax.value, mask

    # -- standalone semi-global block -----------------------------------------------------
    def semi_global_block(self, block, x):
        """SemiGlobalBlock.forward on an fp32 (N, 64, L) CUDA tensor (models/stofnet.py:98-117)."""
        require_cuda(x, "input")
        if x.dim() != 3 or x.shape[1] != 64:
            raise RuntimeError("stofnet_b200: SemiGlobalBlock expects (N, 64, L), got %s" % (tuple(x.shape),))
        dev = x.device
        x = x.detach().to(torch.float32).contiguous()
        N, _, L = x.shape
        with torch.cuda.device(dev):
            # the C handle describes a whole StofNet: the block's own tensors plus zero stand-ins for the convs it never runs
            z = lambda *shape: torch.zeros(shape, dtype=torch.float32)
            sd = {"conv1.weight": z(64, 1, 9), "conv1.bias": z(64), "conv_last.weight": z(4, 64, 3), "conv_last.bias": z(4)}
            for i in (2, 3, 4):
                sd["conv%d.weight" % i], sd["conv%d.bias" % i] = z(64, 64, 7), z(64)
            for k, v in block.state_dict(keep_vars=True).items():
                sd["semi_global_block." + k] = v
            key = (dev.index, tuple((k, v.data_ptr(), v._version) for k, v in block.state_dict(keep_vars=True).items()))
            if self.handle is None or key != self.key:
                W, keep = self._pack(sd)
                W.sgb_scale = int(block.sample_scale)
                self.close()
                h = C.c_void_p()
                check(lib.stof_model_create(C.byref(W), dev.index if dev.index is not None else 0, C.byref(h)))
                self.handle, self.device, self.key = h, dev, key
                del keep
            need = C.c_size_t()
            check(lib.stof_sgb_workspace_bytes(self.handle, N, L, C.byref(need)))
            ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
            out = torch.empty_like(x)
            check(lib.stof_semi_global_block(self.handle, x.data_ptr(), N, L, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                             stream_ptr(dev)))
        return out
