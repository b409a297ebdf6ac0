from .raymarching import (near_far_from_aabb, sph_from_ray, morton3D, morton3D_invert, packbits, march_rays_train,  # noqa: F401
                          composite_rays_train, march_rays, composite_rays, compact_alive, mark_untrained_grid, _backend)
