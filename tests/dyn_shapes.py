"""local_sdf programs (include/ottomarcher.h, OM_SDF_*) used by the dyn-marched tests: the reference's own three SDFs
(marched.rs:57,86-89,134-138) restated as programs, and two shapes the reference does not ship."""


def torus_program(R, r):  # marched.rs:134-138: q = (|(p.x, p.z, 0)| - R, p.y, 0); |q| - r
    return ["x", "z", "len2", R, "sub", "y", "len2", r, "sub"]


def box_program(sx, sy, sz):  # marched.rs:86-89: q = |p| - sizes; |max(q, 0)| + min(max(q.x, max(q.y, q.z)), 0)
    return ["x", "abs", sx, "sub", ("store", 0), "y", "abs", sy, "sub", ("store", 1), "z", "abs", sz, "sub", ("store", 2),
            ("load", 0), 0.0, "max", ("load", 1), 0.0, "max", ("load", 2), 0.0, "max", "len3",
            ("load", 0), ("load", 1), ("load", 2), "max", "max", 0.0, "min", "add"]


def sphere_program(radius):  # marched.rs:57
    return ["x", "y", "z", "len3", radius, "sub"]


def rounded_box_program(sx, sy, sz, rr):  # a user shape: the box SDF minus a rounding radius
    return box_program(sx, sy, sz) + [rr, "sub"]


def capsule_y_program(h, r):  # a user shape: segment (0,-h,0)-(0,h,0) thickened by r: |(x, y - clamp(y,-h,h), z)| - r
    return ["x", "y", "y", -h, "max", h, "min", "sub", "z", "len3", r, "sub"]


def dyn_from_matrix(api, m_local_to_world, code, material, world_f=1.0, bounds=None):
    """DynMarched whose to_local is the homogeneous inverse of m_local_to_world (rotation + translation)."""
    return api.DynMarched(m_local_to_world.fast_homogenous_inverse(), m_local_to_world, world_f, api.sdf_program(code), material, bounds)


def build_dyn_scene(api, with_bounds=True):
    """A ground sphere, a few traced spheres, one typed marched box and four dyn marched objects (torus and box restated,
    rounded box, scaled capsule)."""
    M, w = api.Material, api.HittableList()
    w += api.Sphere.new_with_radius((0, -1000, 0), 1000.0, M.new_lambertian((0.5, 0.5, 0.5)))
    w += api.Sphere.new_with_radius((-4, 1, 0), 1.0, M.new_lambertian((0.4, 0.2, 0.1)))
    w += api.Sphere.new_with_radius((4, 1, 0), 1.0, M.new_metal_fuzz((0.7, 0.6, 0.5), 0.1))
    w += api.Sphere.new_with_radius((0, 1, 2.2), 0.6, M.new_dielectric(1.5))
    w += api.MarchedBox((2.5, 0.4, 2.0), (0.4, 0.4, 0.4), M.new_lambertian((0.2, 0.8, 0.2)))
    T, RX, RZ = api.Mat4x4.new_translate, api.Mat4x4.new_rotate_x, api.Mat4x4.new_rotate_z

    def b(c, r):
        return ((c[0] - r, c[1] - r, c[2] - r), (c[0] + r, c[1] + r, c[2] + r)) if with_bounds else None
    w += dyn_from_matrix(api, T((0, 1, 0)) ^ RX(0.6) ^ RZ(1.33), torus_program(0.5, 0.1), M.new_dielectric(1.5), bounds=b((0, 1, 0), 0.7))
    w += dyn_from_matrix(api, T((-2, 0.5, 2)) ^ RZ(0.3), box_program(0.5, 0.4, 0.3), M.new_metal((0.9, 0.9, 0.9)), bounds=b((-2, 0.5, 2), 0.8))
    w += dyn_from_matrix(api, T((1.2, 0.45, 3.0)) ^ RX(0.2), rounded_box_program(0.3, 0.3, 0.2, 0.1), M.new_lambertian((0.8, 0.3, 0.3)), bounds=b((1.2, 0.45, 3.0), 0.7))
    # capsule given in a frame twice as large as the world's: to_local scales by 2, to_world_f by 0.5
    m = T((-0.5, 0.7, 3.2)) ^ RZ(0.9)
    w2l = api.Mat4x4.new_scale((2, 2, 2)) ^ m.fast_homogenous_inverse()
    w += api.DynMarched(w2l, m, 0.5, api.sdf_program(capsule_y_program(0.8, 0.3)), M.new_metal_fuzz((0.6, 0.7, 0.9), 0.05), b((-0.5, 0.7, 3.2), 0.8))
    return w
