"""Fixed-degree BVH (API of ``chroma/bvh/simple.py``): Morton-sorted leaves, then every
`degree` consecutive nodes share a parent."""
import numpy as np

from chroma_b200.bvh.bvh import BVH
from chroma_b200.gpu.bvh import create_leaf_nodes, merge_nodes, concatenate_layers


def make_simple_bvh(mesh, degree):
    world_coords, leaf_nodes, morton_codes = create_leaf_nodes(mesh)
    order = np.argsort(morton_codes, kind='stable')
    layers = [leaf_nodes[order]]
    while len(layers[0]) > 1:
        layers = [merge_nodes(layers[0], degree)] + layers
    nodes, layer_bounds = concatenate_layers(layers)
    return BVH(world_coords, nodes, [int(x) for x in layer_bounds[:-1]])
