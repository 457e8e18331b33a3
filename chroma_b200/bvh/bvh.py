"""BVH container and node packing (API of ``chroma/bvh/bvh.py``).

Node = uint4: x, y, z hold the lower bound in the low 16 bits and the upper bound in the
high 16 bits (fixed point, see WorldCoords); w holds the child index in the low 28 bits and
the number of children in the high 4 bits (0 = leaf, child is then the triangle id).
"""
import numpy as np

from chroma_b200.gpu.tools import uint4  # noqa: F401  (re-exported like the reference)

CHILD_BITS = 28
NCHILD_MASK = np.uint32(0xFFFF << CHILD_BITS & 0xFFFFFFFF)


def unpack_nodes(nodes):
    "Record array (xlo,xhi,ylo,yhi,zlo,zhi uint16; child uint32; nchild uint16) of packed nodes."
    dtype = np.dtype([('xlo', np.uint16), ('xhi', np.uint16), ('ylo', np.uint16),
                      ('yhi', np.uint16), ('zlo', np.uint16), ('zhi', np.uint16),
                      ('child', np.uint32), ('nchild', np.uint16)])
    out = np.empty(shape=len(nodes), dtype=dtype)
    for axis in 'xyz':
        out[axis + 'lo'] = nodes[axis] & 0xFFFF
        out[axis + 'hi'] = nodes[axis] >> 16
    out['child'] = nodes['w'] & ~NCHILD_MASK
    out['nchild'] = nodes['w'] >> CHILD_BITS
    return out


class OutOfRangeError(Exception):
    "World coordinates do not fit the unsigned 16-bit fixed-point range."


class WorldCoords(object):
    "world = fixed * world_scale + world_origin"
    MAX_INT = 2 ** 16 - 1

    def __init__(self, world_origin, world_scale):
        self.world_origin = np.array(world_origin, dtype=np.float32)
        self.world_scale = np.float32(world_scale)

    def world_to_fixed(self, world):
        "Round world coordinates to the nearest fixed-point value (uint16)."
        fixed = ((np.asarray(world, dtype=float) - self.world_origin) / self.world_scale).round()
        if int(fixed.max()) > WorldCoords.MAX_INT or fixed.min() < 0:
            raise OutOfRangeError('range = (%f, %f)' % (fixed.min(), fixed.max()))
        return fixed.astype(np.uint16)

    def fixed_to_world(self, fixed):
        return np.asarray(fixed) * self.world_scale + self.world_origin


class BVH(object):
    """All nodes in one array, root first, layers contiguous and in order of depth.
    ``layer_offsets[i]`` is the index of the first node of layer i."""

    def __init__(self, world_coords, nodes, layer_offsets):
        self.world_coords = world_coords
        self.nodes = nodes
        self.layer_offsets = layer_offsets
        self.layer_bounds = list(layer_offsets) + [len(nodes)]

    def get_layer(self, layer_number):
        s = slice(self.layer_bounds[layer_number], self.layer_bounds[layer_number + 1])
        return BVHLayerSlice(world_coords=self.world_coords, nodes=self.nodes[s])

    def layer_count(self):
        return len(self.layer_offsets)

    def __len__(self):
        return len(self.nodes)


def node_areas(nodes):
    "Surface area of each node's box in fixed-point units."
    u = unpack_nodes(nodes)
    dx = u['xhi'].astype(float) - u['xlo']
    dy = u['yhi'].astype(float) - u['ylo']
    dz = u['zhi'].astype(float) - u['zlo']
    return 2.0 * (dx * dy + dy * dz + dz * dx)


class BVHLayerSlice(object):
    "One layer of a BVH; modifying ``nodes`` modifies the parent array."

    def __init__(self, world_coords, nodes):
        self.world_coords = world_coords
        self.nodes = nodes

    def __len__(self):
        return len(self.nodes)

    def areas_fixed(self):
        return node_areas(self.nodes)

    def area_fixed(self):
        return node_areas(self.nodes).sum()

    def area(self):
        return self.area_fixed().sum() * self.world_coords.world_scale ** 2

    def get_bounds(self):
        u = unpack_nodes(self.nodes)
        lo = np.dstack([u[s] for s in ('xlo', 'ylo', 'zlo')]).squeeze()
        hi = np.dstack([u[s] for s in ('xhi', 'yhi', 'zhi')]).squeeze()
        return (np.atleast_2d(self.world_coords.fixed_to_world(lo)),
                np.atleast_2d(self.world_coords.fixed_to_world(hi)))
