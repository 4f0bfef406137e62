"""ggnn_b200 -- B200-native GGNN propagation (drop-in for crismolav/ggnn's
``compute_final_node_representations``).  CUDA only; importing the package does not load
the shared library, calling any op does (and raises if it is not built)."""
from .weights import GGNNWeights, load_checkpoint, save_checkpoint, as_weights, from_dict  # noqa: F401
from .engine import Propagator  # noqa: F401
from .api import (compute_final_node_representations,  # noqa: F401
                  compute_final_node_representations_sparse, compact_adjacency, get_propagator)

__all__ = ["GGNNWeights", "load_checkpoint", "save_checkpoint", "as_weights", "from_dict",
           "Propagator", "compute_final_node_representations",
           "compute_final_node_representations_sparse", "compact_adjacency", "get_propagator"]
