"""chroma_b200.bvh: Bounding Volume Hierarchy generation and manipulation."""
from chroma_b200.bvh.bvh import *  # noqa: F401,F403
from chroma_b200.bvh.bvh import BVH, WorldCoords, unpack_nodes, node_areas, BVHLayerSlice, \
    OutOfRangeError, CHILD_BITS, NCHILD_MASK
from chroma_b200.bvh.grid import make_recursive_grid_bvh
from chroma_b200.bvh.simple import make_simple_bvh
