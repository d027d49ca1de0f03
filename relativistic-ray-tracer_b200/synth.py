"""Procedural scenes for the configurations the reference ships no file for: the Cornell
box walls and area light of CBspheres_lambertian with a closed, radially displaced UV sphere of about a million
triangles in place of the two spheres.  Host-side plumbing only: numpy in, SceneData out (and optionally a COLLADA
file the reference's loader accepts, so that the CPU reference can be timed on the same geometry).

The geometry is deterministic in (n_u, n_v, seed): value noise on a 3-D lattice filled by numpy's PCG64 with that seed.
"""
import numpy as np

from . import scene as _sc


def _value_noise(p, seed, cells=16):
    """trilinear value noise with smoothstep weights; p: [n, 3] points in [0, cells)"""
    rng = np.random.default_rng(seed)
    lat = rng.random((cells + 1, cells + 1, cells + 1))
    lat[-1] = lat[0]; lat[:, -1] = lat[:, 0]; lat[:, :, -1] = lat[:, :, 0]      # periodic
    i = np.floor(p).astype(np.int64)
    f = p - i
    i %= cells
    w = f * f * (3.0 - 2.0 * f)
    out = np.zeros(len(p))
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                wt = (w[:, 0] if dx else 1 - w[:, 0]) * (w[:, 1] if dy else 1 - w[:, 1]) * (w[:, 2] if dz else 1 - w[:, 2])
                out += wt * lat[i[:, 0] + dx, i[:, 1] + dy, i[:, 2] + dz]
    return out


def displaced_sphere(n_u=708, n_v=709, centre=(0.0, 0.5, 0.0), radius=0.45, amplitude=0.05, seed=184, frequency=6.0):
    """Closed UV sphere: n_u segments around, n_v stacks, pole fans at both ends: 2 n_u (n_v - 1) triangles
    (1,002,528 for the defaults).  Returns (vertices [nv, 3], faces [nt, 3] int64, counter-clockwise seen from outside)."""
    if n_u < 3 or n_v < 2:
        raise ValueError("displaced_sphere: need n_u >= 3 and n_v >= 2")
    theta = np.pi * np.arange(1, n_v) / n_v                      # interior rings
    phi = 2.0 * np.pi * np.arange(n_u) / n_u
    st, ct = np.sin(theta)[:, None], np.cos(theta)[:, None]
    ring = np.stack([st * np.cos(phi)[None, :], np.broadcast_to(ct, (n_v - 1, n_u)), st * np.sin(phi)[None, :]], axis=-1).reshape(-1, 3)
    dirs = np.concatenate([[[0.0, 1.0, 0.0]], ring, [[0.0, -1.0, 0.0]]])
    noise = _value_noise((dirs * 0.5 + 0.5) * frequency + 0.25, seed)
    r = radius + amplitude * (2.0 * noise - 1.0)
    verts = np.asarray(centre)[None, :] + dirs * r[:, None]
    top, bottom = 0, len(dirs) - 1
    idx = lambda k, j: 1 + k * n_u + (j % n_u)
    j = np.arange(n_u)
    faces = [np.stack([np.full(n_u, top), idx(0, j + 1), idx(0, j)], axis=1)]
    for k in range(n_v - 2):
        a, b, c, d = idx(k, j), idx(k, j + 1), idx(k + 1, j), idx(k + 1, j + 1)
        faces.append(np.stack([a, b, d], axis=1))
        faces.append(np.stack([a, d, c], axis=1))
    faces.append(np.stack([np.full(n_u, bottom), idx(n_v - 2, j), idx(n_v - 2, j + 1)], axis=1))
    return verts, np.concatenate(faces).astype(np.int64)


def vertex_normals(verts, faces):
    """area-weighted vertex normals (sum of the face cross products around a vertex, normalised): what the reference's
    half-edge mesh computes instead of reading the file's normals (reference src/halfEdgeMesh.h:492-515)"""
    p0, p1, p2 = verts[faces[:, 0]], verts[faces[:, 1]], verts[faces[:, 2]]
    fn = np.cross(p1 - p0, p2 - p0)
    n = np.zeros_like(verts)
    for k in range(3):
        np.add.at(n, faces[:, k], fn)
    ln = np.linalg.norm(n, axis=1)
    ln[ln == 0] = 1.0
    return n / ln[:, None]


def cornell_box_with_mesh(base, verts, faces, material=None):
    """`base`: a SceneData whose triangles are the Cornell box walls + light quad (CBspheres_lambertian); its spheres are
    dropped and the mesh appended after the walls in primitive order."""
    nrm = vertex_normals(verts, faces)
    tri_pos = verts[faces].reshape(-1, 9)
    tri_nrm = nrm[faces].reshape(-1, 9)
    if material is None:
        diffuse = [i for i in range(len(base.mats)) if base.mats[i]["type"] == _sc.MAT_DIFFUSE]
        material = diffuse[-1] if diffuse else 0
    nb = base.tri_pos.shape[0]
    nt = nb + tri_pos.shape[0]
    return _sc.SceneData(np.zeros(nt, np.uint8), np.arange(nt, dtype=np.uint32),
                         np.concatenate([base.tri_pos, tri_pos]), np.concatenate([base.tri_nrm, tri_nrm]),
                         np.concatenate([base.tri_mat, np.full(tri_pos.shape[0], material, np.int32)]),
                         np.zeros((0, 4)), np.zeros(0, np.int32), base.mats, base.lights, base.camera)


def c5_scene(base, n_u=708, n_v=709, seed=184):
    """about one million triangles (nodes + FP64 leaf records > the 126 MB L2)"""
    v, f = displaced_sphere(n_u, n_v, seed=seed)
    return cornell_box_with_mesh(base, v, f)


# ---------------------------------------------------------------- COLLADA export ("written as a .dae so both sides load the same file")
def _fmt(a):
    return " ".join(repr(float(x)) for x in np.asarray(a).ravel())


def write_cornell_dae(path, template_text, verts, faces, to_file_axes=None):
    """Writes a COLLADA file the reference's loader (src/collada/collada.cpp) and host/dae_loader.cpp both accept: the Cornell
    box of `template_text` (the text of CBspheres_lambertian.dae) with its two spheres replaced by one triangle mesh bound to
    the `lambertian` material.  Vertex positions are written as float32 (the loaders parse floats), so the scene that comes
    back is `verts.astype(float32)`, not `verts`.  `to_file_axes(v)` maps world positions to what must stand in the file so
    that the loaders' node transform + up-axis handling return them (default: the Z_UP convention of the bundled scenes)."""
    import re
    v = np.asarray(verts, np.float64)
    if to_file_axes is None:
        to_file_axes = lambda w: np.stack([-w[:, 0], w[:, 2], w[:, 1]], axis=1)      # Z_UP file as the loaders read it: world = (-x, z, y)_file
    vf = to_file_axes(v).astype(np.float32)
    f = np.asarray(faces, np.int64)
    t = re.sub(r'\s*<node id="Sphere[12]".*?</node>', "", template_text, flags=re.S)
    geo = ('    <geometry id="bump-mesh" name="bump">\n      <mesh>\n        <source id="bump-mesh-positions">\n'
           '          <float_array id="bump-mesh-positions-array" count="%d">%s</float_array>\n'
           '          <technique_common>\n            <accessor source="#bump-mesh-positions-array" count="%d" stride="3">\n'
           '              <param name="X" type="float"/>\n              <param name="Y" type="float"/>\n              <param name="Z" type="float"/>\n'
           '            </accessor>\n          </technique_common>\n        </source>\n'
           '        <vertices id="bump-mesh-vertices">\n          <input semantic="POSITION" source="#bump-mesh-positions"/>\n        </vertices>\n'
           '        <polylist material="lambertian" count="%d">\n          <input semantic="VERTEX" source="#bump-mesh-vertices" offset="0"/>\n'
           '          <vcount>%s</vcount>\n          <p>%s</p>\n        </polylist>\n      </mesh>\n    </geometry>\n'
           % (vf.size, " ".join(repr(float(x)) for x in vf.ravel()), len(vf), len(f), "3 " * len(f), " ".join(str(int(i)) for i in f.ravel())))
    node = ('      <node id="bump" name="bump" type="NODE">\n        <matrix sid="transform">1 0 0 0 0 1 0 0 0 0 1 0 0 0 0 1</matrix>\n'
            '        <instance_geometry url="#bump-mesh">\n          <bind_material>\n            <technique_common>\n'
            '              <instance_material symbol="lambertian" target="#lambertian"/>\n            </technique_common>\n'
            '          </bind_material>\n        </instance_geometry>\n      </node>\n')
    t = t.replace("  </library_geometries>", geo + "  </library_geometries>", 1)
    t = t.replace("    </visual_scene>", node + "    </visual_scene>", 1)
    with open(path, "w") as fh:
        fh.write(t)
    return vf


def write_c5_dae(path, template_text, n_u=708, n_v=709, seed=184):
    v, f = displaced_sphere(n_u, n_v, seed=seed)
    return write_cornell_dae(path, template_text, v, f), f
