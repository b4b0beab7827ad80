"""Shared helpers for the parity tests: configs of BASELINE.json, seeded synthetic inputs, oracle glue."""
import numpy as np

from oracle import tcnn_oracle as O

CONFIG_1 = {  # BASELINE.json configs[0]: README.md:19-42 inline example of the reference
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 2.0},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_SHIPPED = {  # data/config_hash.json as shipped in the fork
    "loss": {"otype": "RelativeL2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_NERF = {  # BASELINE.json configs[2]: 3-D, per_level_scale 1.38
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.38},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_F4 = {  # F = 4, small table, 3 hidden layers: exercises the other template branches
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam"},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 14, "base_resolution": 8, "per_level_scale": 1.7},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 3},
}

CONFIG_TILED_SMOOTH = {  # TiledGrid + Smoothstep interpolation (grid.h:702-712, common_device.h:801-807), 2-D
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "TiledGrid", "n_levels": 8, "n_features_per_level": 2, "base_resolution": 16, "per_level_scale": 1.5, "interpolation": "Smoothstep"},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_NEAREST = {  # Nearest "interpolation" (grid.h:121-139, 279-282): hashed and dense levels, 3-D
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 12, "base_resolution": 4, "per_level_scale": 1.6, "interpolation": "Nearest"},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_STOCHASTIC = {  # stochastic_interpolation (grid.h:284-303), Smoothstep weights, 2-D
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 12, "base_resolution": 8, "per_level_scale": 1.5,
                 "interpolation": "Smoothstep", "stochastic_interpolation": True},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
CONFIG_DENSE = {  # DenseGrid, Linear, 3-D, F = 4
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "DenseGrid", "n_levels": 4, "n_features_per_level": 4, "base_resolution": 8, "per_level_scale": 1.6},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}
def _net(n, h, act, out_act):
    return {"otype": "FullyFusedMLP", "activation": act, "output_activation": out_act, "n_neurons": n, "n_hidden_layers": h}


_GRID_SMALL_2D = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 12, "base_resolution": 8, "per_level_scale": 1.5}
_GRID_SMALL_3D = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 12, "base_resolution": 8, "per_level_scale": 1.5}
# hidden / output activations other than ReLU / None (common_device.h:100-304), on both MLP widths
# the options the general (one thread per sample and level) grid kernels exist for: F = 1 (fp32 gradient accumulation, grid.h:655-666),
# the hashes other than CoherentPrime (common_device.h:631-688), the per-sample max level (grid.h:69-75)
CONFIG_F1 = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(64, 2, "ReLU", "None"),
             "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 1, "log2_hashmap_size": 12, "base_resolution": 8, "per_level_scale": 1.5}}
CONFIG_HASH_PRIME = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(64, 2, "ReLU", "None"),
                     "encoding": dict(_GRID_SMALL_3D, hash="Prime")}
CONFIG_HASH_REVERSED = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(64, 2, "ReLU", "None"),
                        "encoding": dict(_GRID_SMALL_2D, hash="ReversedPrime", interpolation="Smoothstep")}
CONFIG_HASH_RNG = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(64, 2, "ReLU", "None"),
                   "encoding": dict(_GRID_SMALL_3D, hash="Rng", n_features_per_level=2)}
CONFIG_SMALL_2D = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(64, 2, "ReLU", "None"), "encoding": _GRID_SMALL_2D}

# the reference's 16- and 32-neuron FullyFusedMLP (fully_fused_mlp.cu:862-870): zero-padded inside the 64-wide kernel here
CONFIG_NARROW32 = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "network": _net(32, 3, "ReLU", "None"), "encoding": _GRID_SMALL_2D}
CONFIG_NARROW16 = {"loss": {"otype": "RelativeL2"}, "optimizer": {"otype": "Adam"}, "network": _net(16, 2, "Sigmoid", "None"), "encoding": _GRID_SMALL_3D}

CONFIG_SNAPSHOT = {  # small enough that a full Trainer::serialize(true) snapshot of the reference is a 250 KB fixture
    "loss": {"otype": "L2"}, "optimizer": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 2, "log2_hashmap_size": 10, "base_resolution": 8, "per_level_scale": 1.5},
    "network": _net(64, 2, "ReLU", "None")}
CONFIG_ACT_A = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "encoding": _GRID_SMALL_2D, "network": _net(64, 3, "Sigmoid", "Exponential")}
CONFIG_ACT_B = {"loss": {"otype": "RelativeL2"}, "optimizer": {"otype": "Adam"}, "encoding": {"otype": "OneBlob", "n_bins": 16}, "network": _net(128, 2, "Tanh", "Sigmoid")}
CONFIG_ACT_C = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "encoding": {"otype": "OneBlob", "n_bins": 16}, "network": _net(64, 2, "Squareplus", "Softplus")}
CONFIG_ACT_D = {"loss": {"otype": "L2"}, "optimizer": {"otype": "Adam"}, "encoding": _GRID_SMALL_3D, "network": _net(128, 2, "LeakyReLU", "LeakyReLU")}
CONFIG_ONEBLOB = {  # BASELINE.json configs[1]: data/config_oneblob.json of the reference (bench_ours.cu), 2-D input -> 128 features
    "loss": {"otype": "RelativeL2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-8, "l2_reg": 1e-8},
    "encoding": {"otype": "OneBlob", "n_bins": 64},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 128, "n_hidden_layers": 5},
}
CONFIG_ONEBLOB_64 = {  # the 64-neuron curve of the same benchmark (bench_ours.cu runs config_oneblob.json at 64 and 128 neurons): 128 -> 64 x 5
    "loss": {"otype": "RelativeL2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-8, "l2_reg": 1e-8},
    "encoding": {"otype": "OneBlob", "n_bins": 64},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 5},
}
CONFIG_ONEBLOB16 = {  # OneBlob in front of the 64-neuron kernel (32 features)
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "OneBlob", "n_bins": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 4},
}
CONFIG_WIDE_F4 = {  # BASELINE.json configs[3] with a table small enough for fixtures: 3-D, 16 x 4 features, 128 x 4
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 4, "log2_hashmap_size": 14, "base_resolution": 16, "per_level_scale": 1.38},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 128, "n_hidden_layers": 4},
}
CONFIG_WIDE_DP = {  # BASELINE.json configs[4] with a small table: 3-D, 16 x 2 features, 128 x 3
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "Adam", "learning_rate": 1e-3},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.38},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 128, "n_hidden_layers": 3},
}


def is_oneblob(cfg):
    return cfg["encoding"].get("otype", "").lower() == "oneblob"


def oracle_configs(cfg, n_in, n_out, scales=None):
    e = cfg["encoding"]
    n = cfg["network"]
    if is_oneblob(cfg):  # no table: g is None, the MLP input is the padded OneBlob width
        m = O.MlpConfig(O.next_multiple(n_in * e.get("n_bins", 16), 16), n_out, n.get("n_neurons", 128), n.get("n_hidden_layers", 5),
                        n.get("activation", "ReLU"), n.get("output_activation", "None"))
        return None, m
    otype = e.get("otype", "Grid").lower()
    gtype = "Dense" if otype == "densegrid" else "Tiled" if otype == "tiledgrid" else e.get("type", "Hash")
    g = O.GridConfig(n_in, e.get("n_levels", 16), e.get("n_features_per_level", 2), e.get("log2_hashmap_size", 19),
                     e.get("base_resolution", 16), e.get("per_level_scale", 2.0), scales=scales, grid_type=gtype,
                     interpolation=e.get("interpolation", "Linear"), stochastic=bool(e.get("stochastic_interpolation", False)),
                     hash=e.get("hash", "CoherentPrime"))
    n = cfg["network"]
    m = O.MlpConfig(O.next_multiple(g.n_features, 16), n_out, n.get("n_neurons", 128), n.get("n_hidden_layers", 5),
                    n.get("activation", "ReLU"), n.get("output_activation", "None"))
    return g, m


def synth_inputs(B, n_in, n_out, seed=0):
    """x ~ U[0,1)^D and the analytic target of oracle/ref_driver.cu synth_target."""
    rng = np.random.default_rng(seed)
    x = rng.random((B, n_in), dtype=np.float32)
    s = (x * (1.0 + 0.5 * np.arange(n_in, dtype=np.float32))).sum(axis=1, dtype=np.float32)
    t = np.stack([0.5 + 0.5 * np.sin(6.2831853 * (c + 1) * s) for c in range(n_out)], axis=1).astype(np.float32)
    return x, t


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def oracle_encode(cfg, g, m, x, table16):
    """fp16 [B, m.n_input]: the encoded batch as the MLP sees it (grid features zero-padded, OneBlob one-padded)."""
    if is_oneblob(cfg):
        return O.oneblob_forward(x, cfg["encoding"].get("n_bins", 16), m.n_input)
    enc = np.zeros((x.shape[0], m.n_input), np.float16)
    enc[:, : g.n_features] = O.grid_forward(g, x, table16)
    return enc


# stand-alone grid encodings of the second-order tests / goldens: name -> (n_in, encoding config)
SECOND_ORDER_ENCODINGS = {
    "hash2d_linear": (2, {"otype": "HashGrid", "n_levels": 6, "n_features_per_level": 2, "log2_hashmap_size": 10, "base_resolution": 4, "per_level_scale": 1.6}),
    "hash3d_smooth": (3, {"otype": "HashGrid", "n_levels": 4, "n_features_per_level": 4, "log2_hashmap_size": 9, "base_resolution": 4, "per_level_scale": 1.5, "interpolation": "Smoothstep"}),
    "tiled2d_smooth": (2, {"otype": "TiledGrid", "n_levels": 4, "n_features_per_level": 2, "base_resolution": 8, "per_level_scale": 1.5, "interpolation": "Smoothstep"}),
}


def second_order_inputs(n_in, n_table, n_features, padded_width, B=256, seed=5):
    """Seeded inputs of one second-order case: positions, fp16 table, fp16 dL/dy [B][padded] (padding zero), v = dL2/d(dL/dx)."""
    rng = np.random.default_rng(seed)
    x = rng.random((B, n_in), dtype=np.float32)
    table = rng.uniform(-1, 1, n_table).astype(np.float16)
    dy = np.zeros((B, padded_width), np.float16)
    dy[:, :n_features] = rng.standard_normal((B, n_features)).astype(np.float16)
    v = rng.standard_normal((B, n_in)).astype(np.float32)
    return x, table, dy, v


# stand-alone grid encodings in fp32 precision (tcnn.Encoding(dtype=torch.float32)): goldens enc32_*.npz
FP32_ENCODINGS = dict(SECOND_ORDER_ENCODINGS, nearest3d=(3, CONFIG_NEAREST["encoding"]))


def fp32_inputs(n_in, n_table, n_features, padded_width, B=256, seed=6):
    rng = np.random.default_rng(seed)
    x = rng.random((B, n_in), dtype=np.float32)
    table = rng.uniform(-1, 1, n_table).astype(np.float32)
    dy = np.zeros((B, padded_width), np.float32)
    dy[:, :n_features] = rng.standard_normal((B, n_features)).astype(np.float32)
    return x, table, dy


# goldens dumped with GridEncoding::set_max_level(value) (REF_MAX_LEVEL of `ref_driver dump`): name -> max_level
MAX_LEVEL_CASES = {"maxlevel_f4_3d_T14": 0.5}
# goldens dumped with GridEncoding::set_max_level_gpu (REF_MAX_LEVEL_GPU): sample i keeps the fraction 0.25 * (1 + i % 4) of the levels
MAX_LEVEL_GPU_CASES = {"maxlevelgpu_f4_3d_T14"}
# goldens with a second backward pass in GradientMode::Accumulate on top of the first (REF_ACCUMULATE): `grads_accumulated_*`
ACCUMULATE_CASES = {"accumulate_2d_T12"}


def max_level_per_sample(B):
    return (0.25 * (1 + np.arange(B) % 4)).astype(np.float32)


def big_case_params(cfg, n_in, n_out, seed=77):
    """Parameters of the full-size golden (tests/golden/cfg1_2d_T19_b18.npz): the reference's initial MLP weights with a table of
    uniform(-0.5, 0.5) values instead of the initial +-1e-4 - hidden activations of order one, so that gradients at batch 2^18 sit
    well above fp16 underflow (from the initial table the reference's own fp16 split-K weight gradients are 94 % zeros)."""
    g, m = oracle_configs(cfg, n_in, n_out)
    p = O.init_params(m, g, 1337, True)
    p[m.n_params:] = np.random.default_rng(seed).uniform(-0.5, 0.5, g.n_params).astype(np.float32)
    return p
