"""tinycudann-compatible module surface of the B200-native hot path (drop-in for `import tinycudann as tcnn`,
bindings/torch/tinycudann/__init__.py:9-11 of the reference)."""
from .modules import Encoding, Module, Network, NetworkWithInputEncoding, NativeModule, free_temporary_memory  # noqa: F401
from .trainer import Trainer, TrainableModel, create_from_config  # noqa: F401
from .parallel import PeerMemory, shard_bounds  # noqa: F401
