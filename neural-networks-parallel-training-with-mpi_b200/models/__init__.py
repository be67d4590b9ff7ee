from .mlp import MLP, FlatMLP, MLPSpec, Segment  # noqa: F401
