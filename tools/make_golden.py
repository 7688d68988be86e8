#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the reference itself (oracle/_ref/libdrt_ref.so = the reference's
own sources compiled host-side). Runs only in the build container; the outputs are committed.

  kat.npz            known-answer vectors of the leaf functions: pcg_hash, randomFloat, randomUnitVec3,
                     randomUnitSphereVec3, Bounds3f::intersect, Intersection, Camera::GetRay
                     (incl. the edge cases SURVEY.md 8c lists: ray in the triangle plane, |det| ~ 1e-6,
                     origin inside the box, direction components equal to 0)
  scene_<name>.npz   per fixture scene: reference BVH (node table, leaf order), primary-hit map
                     (canonical id, t, u, v) and a small accumulated render, with CRC32s of the raw bits
"""
import ctypes
import os
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracles as O  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")

# scene -> (primary W, H, render W, H, frames, sun, bounces)
SCENE_CASES = {
    "cornell_box": (512, 512, 96, 96, 4, False, 2),       # BASELINE.json config 1: 512x512 primary rays
    "room": (240, 135, 120, 68, 4, True, 8),              # config 3 settings at reduced size
    "dense_monkey": (240, 135, 120, 68, 4, False, 4),     # config 2
    "valorant_ghost": (240, 135, 120, 68, 4, True, 2),    # config 4a
    "animemaid": (240, 135, 120, 68, 4, True, 2),         # config 4b (alpha cut-out)
    "mc_transparency_test": (160, 90, 96, 54, 4, True, 2),
    "bvhsplit_test": (160, 90, 96, 54, 2, True, 2),
    "cube": (96, 54, 64, 36, 2, True, 2),
    "sunshadow_test": (160, 90, 96, 54, 4, True, 2),      # the reference's models/test/sunshadowTest.glb UNCHANGED: its albedo is a JPEG (stb decode path)
}


def crc(a):
    return np.uint32(zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF)


def make_kat(R):
    L = R.L
    rng = np.random.RandomState(1234)
    seeds = np.array([0, 1, 2, 3, 12345, 0xFFFFFFFF, 1920 * 1080 - 1, 0x80000000, 747796405, 2891336453]
                     + list(rng.randint(0, 2 ** 32, 54, dtype=np.uint64)), np.uint32)
    pcg = np.array([L.ref_pcg_hash(int(s)) for s in seeds], np.uint32)
    rf, rf_seed, uv3, uv3_seed, us3, us3_seed = [], [], [], [], [], []
    for s in seeds:
        c = ctypes.c_uint(int(s))
        rf.append(L.ref_random_float(ctypes.byref(c)))
        rf_seed.append(c.value)
        c = ctypes.c_uint(int(s))
        o = (ctypes.c_float * 3)()
        L.ref_random_unit_vec3(ctypes.byref(c), o)
        uv3.append(list(o))
        uv3_seed.append(c.value)
        c = ctypes.c_uint(int(s))
        L.ref_random_unit_sphere_vec3(ctypes.byref(c), o)
        us3.append(list(o))
        us3_seed.append(c.value)

    # slab test
    boxes, ros, rds, slab = [], [], [], []
    def add_box(b, o, d):
        b, o, d = np.array(b, np.float32), np.array(o, np.float32), np.array(d, np.float32)
        boxes.append(b); ros.append(o); rds.append(d)
        slab.append(L.ref_bounds_intersect(O.P(b), O.P(o), O.P(d)))
    unit = [-1, -1, -1, 1, 1, 1]
    add_box(unit, [0, 0, 5], [0, 0, -1])          # axis-aligned, components 0 -> inf invDir
    add_box(unit, [0, 0, 0], [0, 1, 0])           # origin inside
    add_box(unit, [0, 0, 5], [0, 0, 1])           # pointing away
    add_box(unit, [1, 0, 5], [0, 0, -1])          # grazing a face (0*inf)
    add_box(unit, [2, 2, 2], [-1, -1, -1])
    add_box([0, 0, 0, 1, 0, 1], [0.5, 1, 0.5], [0, -1, 0])   # flat box
    add_box([0, 0, 0, 1, 0, 1], [0.5, 0, -1], [0, 0, 1])     # ray inside the flat box's plane
    for _ in range(57):
        lo = rng.uniform(-5, 5, 3); hi = lo + rng.uniform(0, 4, 3)
        add_box(list(lo) + list(hi), rng.uniform(-10, 10, 3), rng.normal(size=3))

    # Moller-Trumbore
    tris, tos, tds, hit, tt, tuvw = [], [], [], [], [], []
    def add_tri(t9, o, d):
        t9, o, d = np.array(t9, np.float32), np.array(o, np.float32), np.array(d, np.float32)
        t = ctypes.c_float(); uvw = (ctypes.c_float * 3)()
        h = L.ref_intersection(O.P(t9), O.P(o), O.P(d), ctypes.byref(t), uvw)
        tris.append(t9); tos.append(o); tds.append(d); hit.append(h)
        tt.append(t.value if h else -1.0); tuvw.append(list(uvw) if h else [0, 0, 0])
    T = [0, 0, 0, 1, 0, 0, 0, 1, 0]
    add_tri(T, [0.25, 0.25, 1], [0, 0, -1])
    add_tri(T, [0.25, 0.25, -1], [0, 0, 1])        # back face: two-sided
    add_tri(T, [0.25, 0.25, 1], [1, 0, 0])         # parallel: ray in a plane parallel to the triangle
    add_tri(T, [0.25, 0.25, 0], [1, 0, 0])         # ray IN the triangle plane
    add_tri(T, [0, 0, 1], [0, 0, -1])              # through vertex 0
    add_tri(T, [0.5, 0.5, 1], [0, 0, -1])          # on the hypotenuse (u+v == 1)
    add_tri(T, [0.5, 0, 1], [0, 0, -1])            # on an edge (v == 0)
    add_tri(T, [0.25, 0.25, 1e-7], [0, 0, -1])     # t below epsilon
    add_tri(T, [0.25, 0.25, 1], [0, 0, -1e-6])     # |det| == 1e-6 exactly at the threshold region
    add_tri(T, [0.25, 0.25, 1], [0, 0, -1.1e-6])
    add_tri(T, [0.25, 0.25, 1], [0, 0, -0.9e-6])
    add_tri([0, 0, 0, 0, 0, 0, 0, 0, 0], [0, 0, 1], [0, 0, -1])  # degenerate
    for _ in range(52):
        add_tri(rng.uniform(-2, 2, 9), rng.uniform(-4, 4, 3), rng.normal(size=3) * rng.choice([1.0, 30.0, 100.0]))

    # camera rays
    cam = O.default_camera_bytes()
    cam_defocus = cam.copy(); cam_defocus[5] = 2.0
    cases, rays, seeds_after = [], [], []
    for c, (x, y, W, H, f) in [(cam, (0, 0, 512, 512, 1)), (cam, (511, 511, 512, 512, 1)), (cam, (100, 37, 1920, 1080, 7)),
                               (cam, (1919, 1079, 1920, 1080, 256)), (cam_defocus, (5, 9, 64, 48, 3)),
                               (cam_defocus, (63, 47, 64, 48, 11))]:
        od = (ctypes.c_float * 6)()
        s = L.ref_camera_get_ray(O.P(c), x, y, W, H, f, od)
        cases.append([x, y, W, H, f, 1 if c is cam_defocus else 0]); rays.append(list(od)); seeds_after.append(s)

    np.savez_compressed(
        os.path.join(OUT, "kat.npz"),
        seeds=seeds, pcg=pcg, random_float=np.array(rf, np.float32), random_float_seed=np.array(rf_seed, np.uint32),
        unit_vec3=np.array(uv3, np.float32), unit_vec3_seed=np.array(uv3_seed, np.uint32),
        unit_sphere_vec3=np.array(us3, np.float32), unit_sphere_vec3_seed=np.array(us3_seed, np.uint32),
        slab_box=np.array(boxes, np.float32), slab_o=np.array(ros, np.float32), slab_d=np.array(rds, np.float32),
        slab_t=np.array(slab, np.float32),
        mt_tri=np.array(tris, np.float32), mt_o=np.array(tos, np.float32), mt_d=np.array(tds, np.float32),
        mt_hit=np.array(hit, np.int32), mt_t=np.array(tt, np.float32), mt_uvw=np.array(tuvw, np.float32),
        cam_default=cam, cam_defocus=cam_defocus, cam_cases=np.array(cases, np.int32), cam_rays=np.array(rays, np.float32),
        cam_seed_after=np.array(seeds_after, np.uint32))


def make_scene(R, name, case):
    pw, ph, rw, rh, frames, sun, bounces = case
    h = R.load(O.scene_path(name), bins=8, leaf=20, iterative=False)
    cam = O.default_camera_bytes()
    pid, t, u, v = R.trace_primary(h, cam, pw, ph, 1)
    st = O.settings_bytes(sun=sun, bounces=bounces)
    acc, rays, _ = R.render(h, cam, st, rw, rh, 1, frames)
    nodes = R.nodes(h)
    np.savez_compressed(
        os.path.join(OUT, "scene_%s.npz" % name),
        primary_size=np.array([pw, ph], np.int32), prim_id=pid,
        # t/u/v on a 4x-strided grid (values) + CRC32 of the full-resolution raw bits
        tuv_stride4=np.stack([a.reshape(ph, pw)[::4, ::4] for a in (t, u, v)]),
        crc_prim_id=crc(pid), crc_t=crc(t), crc_u=crc(u), crc_v=crc(v),
        render_size=np.array([rw, rh, frames, int(sun), bounces], np.int32), acc=acc, rays=np.array(rays, np.uint64),
        settings=st, camera=cam,
        ref_nodes=nodes, ref_leaf_order=R.canonical_ids(h),
        tri_crc=crc(R.triangles_load_order(h)), mat_crc=crc(R.materials(h)),
        tex_crc=np.array([crc(x) for x in R.textures(h)], np.uint32))
    R.free(h)
    print(name, "hits", float((pid >= 0).mean()), "rays", rays, "nodes", nodes.shape[0])


def main():
    R = O.RefOracle()
    only = sys.argv[1:]  # scene names: regenerate just these (the KATs stay as they are)
    if not only:
        make_kat(R)
    for name, case in SCENE_CASES.items():
        if not only or name in only:
            make_scene(R, name, case)


if __name__ == "__main__":
    main()
