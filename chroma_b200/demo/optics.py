"""Materials and surfaces of the demo detector (names of ``chroma/demo/optics.py``)."""
import numpy as np

from chroma_b200.geometry import Material, Surface
from chroma_b200.demo import optics_tables as T


def _prop(obj, name, rows):
    setattr(obj, name, np.array(rows, dtype=np.float32))


vacuum = Material('vacuum')
vacuum.set('refractive_index', 1.0)
vacuum.set('absorption_length', 1e6)
vacuum.set('scattering_length', 1e6)

lambertian_surface = Surface('lambertian_surface')
lambertian_surface.set('reflect_diffuse', 1)

black_surface = Surface('black_surface')
black_surface.set('absorb', 1)

shiny_surface = Surface('shiny_surface')
shiny_surface.set('reflect_specular', 1)

glossy_surface = Surface('glossy_surface')
glossy_surface.set('reflect_diffuse', 0.5)
glossy_surface.set('reflect_specular', 0.5)

# photocathode of the Hamamatsu R7081HQE: equal detect and absorb, rest diffusely reflected
r7081hqe_photocathode = Surface('r7081hqe_photocathode')
_prop(r7081hqe_photocathode, 'detect', T.R7081HQE_DETECT)
_prop(r7081hqe_photocathode, 'absorb', T.R7081HQE_ABSORB)
_prop(r7081hqe_photocathode, 'reflect_diffuse', T.R7081HQE_REFLECT_DIFFUSE)

glass = Material('glass')
_prop(glass, 'refractive_index', T.GLASS_REFRACTIVE_INDEX)
_prop(glass, 'absorption_length', T.GLASS_ABSORPTION_LENGTH)
_prop(glass, 'scattering_length', T.GLASS_SCATTERING_LENGTH)

water = Material('water')
water.density = 1.0
water.composition = {'H': 0.1119, 'O': 0.8881}
_prop(water, 'refractive_index', T.WATER_REFRACTIVE_INDEX)
_prop(water, 'absorption_length', T.WATER_ABSORPTION_LENGTH)
_prop(water, 'scattering_length', T.WATER_SCATTERING_LENGTH)
