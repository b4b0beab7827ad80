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


def oracle_configs(cfg, n_in, n_out, scales=None):
    e = cfg["encoding"]
    g = O.GridConfig(n_in, e.get("n_levels", 16), e.get("n_features_per_level", 2), e.get("log2_hashmap_size", 19),
                     e.get("base_resolution", 16), e.get("per_level_scale", 2.0), scales=scales)
    n = cfg["network"]
    m = O.MlpConfig(O.next_multiple(g.n_features, 16), n_out, n.get("n_neurons", 128), n.get("n_hidden_layers", 5))
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
