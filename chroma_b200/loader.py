"""Geometry assembly helper (the compute part of ``chroma/loader.py``: flatten + BVH).
The reference's on-disk pickle cache (chroma/cache.py) is not part of the engine."""
from chroma_b200.bvh import make_recursive_grid_bvh


def build(geometry, target_degree=3, fetch_nodes=True):
    "Flatten `geometry` and attach a recursive-grid BVH (needs a current context)."
    geometry.flatten()
    geometry.bvh = make_recursive_grid_bvh(geometry.mesh, target_degree=target_degree,
                                           fetch_nodes=fetch_nodes)
    return geometry


def load_bvh(geometry, bvh_name='default', auto_build_bvh=True, read_bvh_cache=True,
             update_bvh_cache=True, cache_dir=None, cuda_device=None, target_degree=3):
    "Signature of chroma/loader.py:131-160; always builds (no cache)."
    return make_recursive_grid_bvh(geometry.mesh, target_degree=target_degree)


def create_geometry_from_obj(obj, bvh_name='default', auto_build_bvh=True, read_bvh_cache=True,
                             update_bvh_cache=True, cache_dir=None, cuda_device=None):
    "Flatten a Geometry/Detector and give it a BVH (chroma/loader.py:162-189)."
    return build(obj)
