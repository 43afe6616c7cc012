"""Shared implementation of the three decoder wrappers.  The public classes
(parallel_decoder.py, sequential_decoder.py, sequential_decoder_reverse.py) only fix the
variant, the chaining order and the scale offset, like their reference counterparts."""
import torch
from torch import nn

from ..decoder_utils.pos_encoding import Embedder
from ..easydict import EasyDict
from . import fused
from .base_decoder import Decoder


class FusedDecoderBase(nn.Module):
    VARIANT = None          # "parallel" | "sequential" | "sequential_reverse"
    ORDER = None            # chaining order of the heads
    SCALE_BIAS = None
    MAPPING_GETS_CAMERA = False   # reference quirk q2 (SequentialDecoder swaps the camera argument)

    def __init__(self, G, hidden_dim=128, use_xyz_embedding=True, use_gen_finetune=True, device="cuda"):
        super().__init__()
        if not 1 <= int(hidden_dim) <= fused.KERNEL_HIDDEN:
            raise NotImplementedError(f"hidden_dim={hidden_dim}: the fused sm_100a kernels are {fused.KERNEL_HIDDEN} units wide; narrower "
                                      "decoders run zero-padded on them, wider ones are not supported")
        self.hidden_dim = hidden_dim
        self.use_xyz_embedding = use_xyz_embedding
        self.use_gen_finetune = use_gen_finetune
        self.device = device
        position_dim = 3
        if use_xyz_embedding:
            self.embedder = Embedder(include_input=True, input_dims=3, num_freqs=10)
            position_dim = self.embedder.out_dim
        self.G = G
        out = {"xyz": 3, "scale": 3, "rotation": 4, "opacity": 1, "color": 3}
        chained = 0
        heads = {}
        for h in self.ORDER:
            extra = 0 if self.VARIANT == "parallel" else chained
            heads[h] = Decoder(n_features=32 + position_dim + extra, out_features=out[h], hidden_dim=hidden_dim)
            chained += out[h]
        # registration order follows the reference constructors (state_dict / parameters() order)
        for h in self.ORDER:
            setattr(self, f"{h}_decoder", heads[h].to(device))
        self.scale_activation = torch.nn.Softplus()
        # reference quirk q1: SequentialDecoderReverse.forward reads self.triplane_sr, which its
        # __init__ never sets; define the only supported value for every variant.
        self.triplane_sr = "None"
        self.precision = fused.DEFAULT_PRECISION     # "fp16" (tensor cores) or "fp32" (exact)
        self._flat = None
        self._cfg_cache = {}
        self._last_activated = None

    # ---- reference API -----------------------------------------------------
    def activate_scale(self, scale):
        return -self.scale_activation(scale + 5) + self.SCALE_BIAS

    def get_params_custom(self):
        params = []
        for h in ("xyz", "scale", "rotation", "opacity", "color"):
            params += list(getattr(self, f"{h}_decoder").parameters())
        if self.use_gen_finetune:
            params += list(self.G.parameters())
        return params

    def forward(self, z, gan_camera_params, init_position, truncation_psi):
        zeros = torch.zeros_like(gan_camera_params)
        map_c, syn_c = (gan_camera_params, zeros) if self.MAPPING_GETS_CAMERA else (zeros, gan_camera_params)
        ws = self.G.mapping(z, map_c, truncation_psi=truncation_psi)
        synth = self.G.synthesis(ws, syn_c, noise_mode="const")
        if self.triplane_sr != "None":
            raise NotImplementedError("triplane super-resolution is not part of the reference's working path")
        return self.decode(synth["feature_planes"], init_position)

    # ---- fused path --------------------------------------------------------
    def _cfg_for(self, planes, precision=None):
        precision = precision or self.precision
        kw = self.G.rendering_kwargs
        depth = int(kw["triplane_depth"]) if "triplane_depth" in kw else 1
        arch = "panohead" if "triplane_depth" in kw else "eg3d"
        key = (arch, depth, planes.shape[-2], planes.shape[-1], float(kw["box_warp"]), precision)
        if key not in self._cfg_cache:
            if planes.shape[2] != 32 * depth:
                raise ValueError(f"feature_planes has {planes.shape[2]} channels, expected 32*{depth}")
            self._cfg_cache[key] = fused.make_config(arch, self.VARIANT, depth, planes.shape[-2], planes.shape[-1],
                                                     10 if self.use_xyz_embedding else 0, fused.KERNEL_HIDDEN,
                                                     kw["box_warp"], self.SCALE_BIAS, precision)
        return self._cfg_cache[key]

    def _ensure_flat(self, planes):
        cfg = self._cfg_for(planes)
        if self._flat is not None and self._flat.cfg.precision != cfg.precision:
            self._flat.set_config(cfg)
        if self._flat is None:
            self._flat = fused.FlatParams(self, cfg)
        elif not self._flat.in_sync(self):
            self._flat.adopt(self)      # parameters were moved / reloaded: re-point them
        return self._flat

    def decode(self, feature_planes, init_position):
        """planes [1,3,32*D,H,W] + positions [P,3] -> EasyDict(xyz, scale, rotation, opacity, color)."""
        flat = self._ensure_flat(feature_planes)
        needs_grad = torch.is_grad_enabled() and (flat.flat.requires_grad or feature_planes.requires_grad or
                                                 any(p.requires_grad for p in self.get_params_custom()[:40]))
        if needs_grad and isinstance(feature_planes, fused.PackedPlanes):
            raise RuntimeError("PackedPlanes are inference-only: run under torch.no_grad() or pass the fp32 planes")
        if needs_grad:
            from . import fused_backward
            xyz, scale, rotation, opacity, color = fused_backward.apply(self, feature_planes, init_position)
        else:
            res = fused.decode_forward(self._cfg_for(feature_planes), feature_planes, init_position,
                                       flat.packed_weights())
            self._last_activated = {k: res[k] for k in ("act_scale", "act_rotation", "act_opacity")}
            xyz, scale, rotation, opacity, color = (res[k] for k in ("xyz", "scale", "rotation", "opacity", "color"))
        result = EasyDict()
        for k in self.ORDER:        # insertion order as in the reference's forward
            result[k] = {"xyz": xyz, "scale": scale, "rotation": rotation, "opacity": opacity, "color": color}[k]
        return result

    def invalidate_packed_weights(self):
        """See ``FlatParams.invalidate``: needed only after parameter writes that bypass autograd's version counters."""
        if self._flat is not None:
            self._flat.invalidate()

    def pack_planes(self, feature_planes):
        """fp32 NCHW planes -> ``fused.PackedPlanes`` (fp16 channels-last), accepted by ``decode`` / ``forward``
        (when the generator returns it as ``feature_planes``) and by ``FramePipeline.frame`` in place of the tensor."""
        return fused.pack_planes(self._cfg_for(feature_planes, precision="fp16"), feature_planes)

    def activated(self):
        """exp(scale), normalize(rotation), sigmoid(opacity) of the LAST decode, produced by the
        same kernel (the GaussianModel getters, gaussian_splatting/scene/gaussian_model.py:99-121)."""
        return self._last_activated
