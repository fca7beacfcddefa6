"""Canonical synthetic model configurations and seeded weights.

The reference loads ``./unified/results/demo_small/{config.yaml,weights.pt}``
(sender/encoder/codec_pipeline.py:56-72); both files live in the git submodule that is empty in the
reference checkout (.SUBMODULES.json:8-15), so channel widths and weights are not recoverable.  This
module freezes the stand-in: a PCGCv2-style sparse auto-encoder with a
mean-scale hyperprior, config driven, with seeded variance-preserving weights.  A real
``state_dict`` with the same key names drops in through ``ColorModel.load_state_dict``.

State-dict layout (numpy float32):
  <layer>.kernel [kernel_volume, C_in, C_out]   (MinkowskiEngine layout)      <layer>.bias [C_out]
  entropy_model.scale_nn.{w1 [2,H], b1 [H], w2 [H,Cy], b2 [Cy]}   entropy_model.offset_gain [Cy]
  entropy_model.entropy_bottleneck.{matrices.i, biases.i, factors.i, quantiles}   (CompressAI shapes)
"""
import math

import numpy as np

SYNTH_SMALL = {
    "name": "synth_small",
    "c_in": 4,              # (1, r, g, b)   codec_pipeline.py:258
    "stem": 32,
    "widths": [32, 64, 128],  # after each stride-2 stage of g_a (tensor stride 2, 4, 8)
    "blocks": 2,            # residual blocks (two k3 convs each) per stage
    "c_y": 64,
    "c_z": 64,
    "c_h": 64,              # hyper transform width
    "scale_nn_hidden": 64,
    "eps": 1e-4,
    "eb_filters": [3, 3, 3, 3],
    "eb_init_scale": 10.0,
}

SYNTH_WIDE = dict(SYNTH_SMALL, name="synth_wide", stem=128, widths=[128, 128, 128], c_y=128, c_z=128, c_h=128)

CONFIGS = {"synth_small": SYNTH_SMALL, "synth_wide": SYNTH_WIDE, "demo_small": SYNTH_SMALL}

DEFAULT_SEED = 1234


def layer_specs(cfg):
    """[(name, kernel_volume, c_in, c_out, effective_fan_offsets, gain)] for every conv layer."""
    L = []
    w = cfg["widths"]
    nb = cfg["blocks"]
    relu = math.sqrt(2.0)
    # ---- g_a
    L.append(("g_a.stem", 27, cfg["c_in"], cfg["stem"], 9.0, relu))
    cprev = cfg["stem"]
    for i, c in enumerate(w):
        L.append((f"g_a.down{i}", 8, cprev, c, 4.0, relu))
        for b in range(nb):
            L.append((f"g_a.block{i}_{b}.conv0", 27, c, c, 9.0, relu))
            L.append((f"g_a.block{i}_{b}.conv1", 27, c, c, 9.0, 0.5))
        cprev = c
    L.append(("g_a.head", 1, cprev, cfg["c_y"], 1.0, 1.0))
    # ---- g_s (mirror)
    L.append(("g_s.pre", 1, cfg["c_y"], w[-1], 1.0, relu))
    ups = [w[1], w[0], cfg["stem"]]
    cprev = w[-1]
    for i, c in enumerate(ups):
        L.append((f"g_s.up{i}", 8, cprev, c, 1.0, relu))
        for b in range(nb):
            L.append((f"g_s.block{i}_{b}.conv0", 27, c, c, 14.0, relu))
            L.append((f"g_s.block{i}_{b}.conv1", 27, c, c, 14.0, 0.5))
        L.append((f"g_s.occ{i}", 27, c, 1, 14.0, 1.0))
        cprev = c
    L.append(("g_s.color", 27, cprev, 3, 9.0, 0.1))
    L.append(("g_s.down_conv", 8, 1, 1, 4.0, 1.0))
    # ---- hyper transforms
    ch = cfg["c_h"]
    L.append(("entropy_model.h_a.down0", 8, cfg["c_y"], ch, 4.0, 0.5))
    L.append(("entropy_model.h_a.down1", 8, ch, cfg["c_z"], 4.0, 4.0))
    L.append(("entropy_model.h_s.up0", 8, cfg["c_z"], ch, 1.0, 0.5))
    L.append(("entropy_model.h_s.up1", 8, ch, ch, 1.0, relu))
    L.append(("entropy_model.h_s.head", 1, ch, 2 * cfg["c_y"], 1.0, 1.0))
    return L


def make_state_dict(cfg=SYNTH_SMALL, seed=DEFAULT_SEED):
    rng = np.random.default_rng(seed)
    sd = {}
    for name, kvol, cin, cout, fan, gain in layer_specs(cfg):
        std = gain / math.sqrt(cin * fan)
        sd[name + ".kernel"] = (rng.standard_normal((kvol, cin, cout)) * std).astype(np.float32)
        sd[name + ".bias"] = (rng.standard_normal(cout) * 0.05).astype(np.float32)
    cy = cfg["c_y"]
    # colour head: centre colours in [0,1]
    sd["g_s.color.bias"] = np.full(3, 0.5, dtype=np.float32)
    # h_s head: first c_y channels are scales (kept mostly positive), last c_y are means
    hb = sd["entropy_model.h_s.head.bias"]
    hb[:cy] = 2.5
    hb[cy:] = 0.0
    hk = sd["entropy_model.h_s.head.kernel"]
    hk[:, :, :cy] *= 0.5
    hk[:, :, cy:] *= 0.3
    # scale_nn: MLP 2 -> H -> c_y, softplus output around 1
    H = cfg["scale_nn_hidden"]
    sd["entropy_model.scale_nn.w1"] = (rng.standard_normal((2, H)) * 0.7).astype(np.float32)
    sd["entropy_model.scale_nn.b1"] = (rng.standard_normal(H) * 0.1).astype(np.float32)
    sd["entropy_model.scale_nn.w2"] = (rng.standard_normal((H, cy)) * 0.08).astype(np.float32)
    sd["entropy_model.scale_nn.b2"] = np.full(cy, 0.2, dtype=np.float32)
    sd["entropy_model.offset_gain"] = np.ones(cy, dtype=np.float32)
    # EntropyBottleneck (CompressAI init)
    filters = [1] + list(cfg["eb_filters"]) + [1]
    scale = cfg["eb_init_scale"] ** (1.0 / (len(cfg["eb_filters"]) + 1))
    cz = cfg["c_z"]
    for i in range(len(filters) - 1):
        init = math.log(math.expm1(1.0 / scale / filters[i + 1]))
        sd[f"entropy_model.entropy_bottleneck.matrices.{i}"] = np.full((cz, filters[i + 1], filters[i]), init,
                                                                      dtype=np.float32)
        sd[f"entropy_model.entropy_bottleneck.biases.{i}"] = rng.uniform(-0.5, 0.5, (cz, filters[i + 1], 1)).astype(
            np.float32)
        if i < len(filters) - 2:
            sd[f"entropy_model.entropy_bottleneck.factors.{i}"] = np.zeros((cz, filters[i + 1], 1), dtype=np.float32)
    q = np.tile(np.array([-cfg["eb_init_scale"], 0.0, cfg["eb_init_scale"]], dtype=np.float32), (cz, 1, 1))
    sd["entropy_model.entropy_bottleneck.quantiles"] = q
    return sd
