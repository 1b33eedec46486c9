"""CPU-side tests of the product: the C ABI surface (symbols, struct layouts, C-cleanliness of the headers),
the host mirror of the reference's scene API and its flattener.  No compute entry point is called (no GPU)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


def _declared_functions(path):
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"^\s*(?:const\s+)?(?:int|void|uint32_t|uint64_t|char\s*\*|RtContext\s*\*)\s*\*?\s*(rt[h]?_[a-z0-9_]+)\s*\(", src, flags=re.M)


def test_library_exports_every_declared_symbol(rt):
    declared = _declared_functions(os.path.join(INC, "rt_b200.h")) + _declared_functions(os.path.join(INC, "rt_b200_host.h"))
    assert len(declared) >= 17
    assert sorted(set(declared)) == sorted(set(rt.ABI_SYMBOLS))
    for name in declared:
        assert hasattr(rt.lib, name), f"librt_b200.so does not export {name}"
    assert rt.lib.rt_abi_version() == 2


def test_headers_are_plain_c_and_layouts_match_ctypes(rt, tmp_path):
    prog = tmp_path / "layout.c"
    prog.write_text(r'''
#include <stdio.h>
#include "rt_b200.h"
#include "rt_b200_host.h"
int main(void) {
    printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(RtPrim), sizeof(RtMaterial), sizeof(RtTexture), sizeof(RtSceneDesc),
           sizeof(RtCamera), sizeof(RtRenderParams), sizeof(RtStats), sizeof(RtRayIn), sizeof(RtHitOut), sizeof(RthCameraArgs),
           sizeof(RtContextOptions), sizeof(RtSceneOptions));
    return 0;
}''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", INC, str(prog), "-o", str(exe)])
    sizes = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    expect = [rt.PRIM_DTYPE.itemsize, rt.MATERIAL_DTYPE.itemsize, rt.TEXTURE_DTYPE.itemsize, C.sizeof(rt.RtSceneDesc), C.sizeof(rt.RtCamera),
              C.sizeof(rt.RtRenderParams), C.sizeof(rt.RtStats), rt.RAY_DTYPE.itemsize, rt.HIT_DTYPE.itemsize, C.sizeof(rt.RthCameraArgs),
              C.sizeof(rt.RtContextOptions), C.sizeof(rt.RtSceneOptions)]
    assert sizes == expect
    assert sizes[0] == 128 and sizes[1] == 16 and sizes[4] == 64 and sizes[10] == 32 and sizes[11] == 32


def test_product_never_touches_the_oracle():
    """The shipped path must not import, link or execute anything under oracle/ (and has no CPU fallback)."""
    pkg = os.path.join(ROOT, "ray-tracer_b200")
    for base, _, files in os.walk(pkg):
        if os.path.basename(base) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="ignore").read()
                code = [ln for ln in text.splitlines() if re.search(r"#\s*include|\bimport\b|CDLL|dlopen|-l\w|\.so\b", ln)]
                assert not any("oracle" in ln for ln in code), f"{f} pulls in the oracle"
    ldd = subprocess.check_output(["ldd", os.path.join(pkg, "librt_b200.so")]).decode()
    assert "oracle" not in ldd


def test_missing_device_fails_loudly(rt):
    try:
        import torch

        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    with pytest.raises(rt.BackendError) as e:
        rt.Device(0)
    assert e.value.code == rt.RT_ERR_NO_DEVICE and "no CPU fallback" in str(e.value)


def test_unknown_scene_and_null_arguments(rt):
    with pytest.raises(rt.BackendError):
        rt.FlatScene("no-such-scene")
    assert rt.lib.rt_render(None, None, None, None, None, None) == rt.RT_ERR_INVALID
    assert rt.lib.rt_trace(None, None, 5, 0, None) == rt.RT_ERR_INVALID
    assert rt.lib.rt_scene_create(None, None, None) == rt.RT_ERR_INVALID
    assert b"null" in rt.lib.rt_last_error()


# ---- flattener vs the oracle's own DFS enumeration of the same scene graph ----------------------------
@pytest.mark.parametrize("name", ["book-one", "cornell-box", "book-two", "stress:300"])
def test_flatten_matches_scene_graph(rt, orc, name):
    flat = rt.FlatScene(name, 1)
    prims, mats, texs = flat.prims(), flat.materials(), flat.textures()
    ref = orc.Scene(name, 1).prims()
    assert len(prims) == len(ref)
    assert list(prims["order"]) == list(range(len(ref)))  # DFS rank = tie-break id (optimize.rs:475,484)
    kind_map = {0: rt.PRIM_SPHERE, 1: rt.PRIM_RECT, 2: rt.PRIM_MEDIUM_SPHERE}
    assert [kind_map[k] for k in ref["kind"]] == list(prims["kind"])
    # composed transforms: oracle is column-major 4x4 f64, the flat record row-major 3x4 f32
    wfl = ref["world_from_local"].reshape(-1, 4, 4).transpose(0, 2, 1)[:, :3, :]
    got = prims["world_from_local"].reshape(-1, 3, 4)
    assert np.allclose(got, wfl, rtol=1e-6, atol=1e-6)
    inv = np.linalg.inv(ref["world_from_local"].reshape(-1, 4, 4).transpose(0, 2, 1))[:, :3, :]
    assert np.allclose(prims["local_from_world"].reshape(-1, 3, 4), inv, rtol=1e-5, atol=1e-4)
    assert np.allclose(prims["param"][:, 0], ref["param"][:, 0], rtol=1e-6)
    # the OUTERMOST sprite's material wins (sprite.rs:124)
    mat_kind = np.where(prims["material"] >= 0, mats["kind"][np.maximum(prims["material"], 0)].astype(np.int64), -1)
    assert list(mat_kind) == list(ref["material_kind"])
    assert np.all((mats["texture"] >= 0) | (mats["kind"] == rt.MAT_DIELECTRIC))
    assert np.all(mats["texture"] < len(texs))


def test_book_one_scene_shape(rt):  # examples/book-one.rs:103-205
    flat = rt.FlatScene("book-one", 1)
    prims, mats = flat.prims(), flat.materials()
    assert 470 <= len(prims) <= 489
    r = prims["param"][:, 0]
    assert sorted(np.unique(r).tolist()) == pytest.approx([0.2, 1.0, 1000.0, 2000.0])
    kinds = mats["kind"][prims["material"]]
    small = r == np.float32(0.2)
    frac = [np.mean(kinds[small] == k) for k in (rt.MAT_LAMBERTIAN, rt.MAT_METAL, rt.MAT_DIELECTRIC)]
    assert abs(frac[0] - 0.3) < 0.08 and abs(frac[1] - 0.3) < 0.08 and abs(frac[2] - 0.4) < 0.08
    assert np.sum(kinds == rt.MAT_DIFFUSE_LIGHT) == 1
    a = flat.camera_args(16 / 9)
    assert list(a.eye) == [13.0, 2.0, 3.0] and a.lens_radius == 0.05 and a.focus_distance == 10.0


def test_book_two_texture_baking(rt):  # ImageTexture closure baked at texel centres (main.rs:267-280)
    flat = rt.FlatScene("book-two", 1)
    texs = flat.textures()
    img = texs[texs["kind"] == rt.TEX_IMAGE]
    assert len(img) == 1 and (img["width"][0], img["height"][0]) == (1024, 512)
    d = flat.desc()
    assert d.n_texels == 1024 * 512
    texels = np.frombuffer((C.c_float * (4 * d.n_texels)).from_address(d.texels), dtype=np.float32).reshape(512, 1024, 4)
    assert texels[..., :3].min() >= 0.0 and texels[..., :3].max() <= 1.0
    assert np.all(texels[..., :3] * 255 == np.round(texels[..., :3] * 255))  # 8-bit source values / 255
    kinds = flat.prims()["kind"]
    assert np.sum(kinds == rt.PRIM_MEDIUM_SPHERE) == 2 and np.sum(kinds == rt.PRIM_RECT) == 2401 and np.sum(kinds == rt.PRIM_SPHERE) == 1005


def test_camera_constants_match_oracle(rt, orc):  # camera.rs:34-45 in f64, rounded to f32
    flat = rt.FlatScene("book-one", 1)
    cam = flat.camera(400 / 225)
    ll, hz, vt = orc.camera_frame(orc.Scene("book-one", 1).camera(400 / 225))
    assert np.array_equal(np.array(cam.lower_left[:]), ll.astype(np.float32))
    assert np.array_equal(np.array(cam.horizontal[:]), hz.astype(np.float32))
    assert np.array_equal(np.array(cam.vertical[:]), vt.astype(np.float32))
    assert cam.lens_radius == np.float32(0.05)


# ---- independent numpy brute force over the FLATTENED tables vs the oracle's BVH traversal -------------
def _brute_force(prims, rays):
    """Closest hit by testing every flattened primitive (f64 numpy, written from SURVEY Appendix A):
    pins scene-graph flattening, transform composition and the oracle's BVH recursion against each other."""
    n = len(rays)
    best_t = np.full(n, np.inf)
    best_order = np.full(n, -1, dtype=np.int64)
    o, d = rays[:, :3], rays[:, 3:]
    for p in prims:
        if p["kind"] == 2:
            continue  # media are stochastic; tested separately
        m = p["local_from_world"].reshape(3, 4).astype(np.float64)
        lo = o @ m[:, :3].T + m[:, 3]
        ld = d @ m[:, :3].T
        if p["kind"] == 0:
            r = float(p["param"][0])
            a = np.sum(ld * ld, axis=1)
            b = 2 * np.sum(lo * ld, axis=1)
            c = np.sum(lo * lo, axis=1) - r * r
            disc = b * b - 4 * a * c
            ok = disc >= 0
            sq = np.sqrt(np.where(ok, disc, 0))
            t1, t2 = (-b - sq) / (2 * a), (-b + sq) / (2 * a)
            t = np.where(t1 > 1e-6, t1, np.where(t2 > 1e-6, t2, np.inf))
            t = np.where(ok, t, np.inf)
        else:
            w, h = float(p["param"][0]), float(p["param"][1])
            with np.errstate(divide="ignore", invalid="ignore"):
                t = -lo[:, 2] / ld[:, 2]
            x, y = lo[:, 0] + ld[:, 0] * t, lo[:, 1] + ld[:, 1] * t
            bad = ~np.isfinite(t) | (t < 1e-6) | (x < -w / 2) | (x > w / 2) | (y < -h / 2) | (y > h / 2)
            t = np.where(bad, np.inf, t)
        win = t < best_t
        best_t = np.where(win, t, best_t)
        best_order = np.where(win, p["order"], best_order)
    return best_t, best_order


@pytest.mark.parametrize("name", ["book-one", "cornell-box"])
def test_oracle_bvh_agrees_with_bruteforce_over_flat_tables(rt, orc, name):
    prims = rt.FlatScene(name, 1).prims()
    s = orc.Scene(name, 1)
    rays = orc.camera_rays(s.camera(1.5), 30, 20)
    hits = s.trace(rays)
    pts = hits["p"][hits["hit"] == 1]
    rng = np.random.default_rng(2)
    dirs = rng.normal(size=(len(pts), 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    bounce = np.concatenate([pts + 1e-3 * dirs * np.maximum(1.0, np.abs(pts).max(axis=1, keepdims=True)), dirs], axis=1)
    rays = np.concatenate([rays, bounce])
    hits = s.trace(rays)
    t, order = _brute_force(prims, rays)
    st = hits["stable"] == 1
    assert st.mean() > 0.9
    assert np.array_equal(np.isfinite(t)[st], hits["hit"][st] == 1)
    assert np.array_equal(order[st], hits["prim"][st])
    both = st & np.isfinite(t)
    tol = 1e-4 * np.maximum(hits["t"][both], 1e-3) + 64 * hits["t_sens"][both]  # f32 matrices in the flat tables
    assert np.all(np.abs(t[both] - hits["t"][both]) <= tol)


def test_image_encoders_match_the_example_drivers_formats(rt):
    """rth_encode_image: PPM P3 text exactly as examples/book-one.rs:28-30,90-100 prints it, and a PNG that
    decodes (zlib + chunk CRCs checked here, PIL when present) to the RGBA pixels examples/main.rs:105-135 puts."""
    import struct
    import zlib

    rng = np.random.default_rng(5)
    for w, h in [(1, 1), (7, 3), (300, 250)]:  # 300x250 RGBA rows exceed one 65535-byte stored block
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        img[0, 0] = (0, 255, 7)
        ppm = rt.encode_image(img, rt.IMAGE_PPM_P3).decode()
        expect = f"P3\n{w} {h}\n255\n" + "".join(f"{int(p[0])} {int(p[1])} {int(p[2])}\n" for p in img.reshape(-1, 3))  # println!("{:?} {:?} {:?}")
        assert ppm == expect
        png = rt.encode_image(img, rt.IMAGE_PNG_RGBA8)
        assert png[:8] == b"\x89PNG\r\n\x1a\n"
        pos, chunks = 8, []
        while pos < len(png):
            (n,) = struct.unpack(">I", png[pos:pos + 4])
            kind, data = png[pos + 4:pos + 8], png[pos + 8:pos + 8 + n]
            (crc,) = struct.unpack(">I", png[pos + 8 + n:pos + 12 + n])
            assert crc == zlib.crc32(kind + data)
            chunks.append((kind, data))
            pos += 12 + n
        assert [k for k, _ in chunks] == [b"IHDR", b"IDAT", b"IEND"]
        assert struct.unpack(">IIBBBBB", chunks[0][1]) == (w, h, 8, 6, 0, 0, 0)
        raw = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(h, 1 + 4 * w)
        assert np.all(raw[:, 0] == 0)
        rgba = raw[:, 1:].reshape(h, w, 4)
        assert np.array_equal(rgba[..., :3], img) and np.all(rgba[..., 3] == 255)
        try:
            import io
            from PIL import Image
            assert np.array_equal(np.asarray(Image.open(io.BytesIO(png)).convert("RGBA"))[..., :3], img)
        except ImportError:
            pass
    with pytest.raises(rt.BackendError):
        rt.encode_image(np.zeros((2, 2, 3), dtype=np.uint8), 9)


def test_image_decoders_and_earth_texture_ingestion(rt):
    """rth_decode_image: PNGs written by an independent encoder (PIL: deflate-compressed, adaptive filters,
    every 8-bit colour type), our own stored-deflate PNGs and PPM P3 / P6 decode to the exact pixels; malformed and unsupported
    streams are refused.  rth_set_earth_image: book-two bakes the installed image 1:1 into its texel table (examples/main.rs:266-281)."""
    import io
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
    img[5:20, 5:30] = (10, 200, 30)  # flat area + noise: PIL picks different filters per row
    assert np.array_equal(rt.decode_image(rt.encode_image(img, rt.IMAGE_PNG_RGBA8)), img)
    assert np.array_equal(rt.decode_image(rt.encode_image(img, rt.IMAGE_PPM_P3)), img)
    p6 = b"P6\n# a comment\n53 37\n255\n" + img.tobytes()
    assert np.array_equal(rt.decode_image(p6), img)
    try:
        from PIL import Image
    except ImportError:
        Image = None
    if Image is not None:
        for mode in ["RGB", "RGBA", "L", "LA", "P"]:
            pil = Image.fromarray(img).convert(mode)
            buf = io.BytesIO()
            pil.save(buf, format="PNG")
            expect = np.asarray(pil.convert("RGB") if mode != "LA" else pil.convert("L").convert("RGB"))
            assert np.array_equal(rt.decode_image(buf.getvalue()), expect), mode
        buf = io.BytesIO()
        Image.fromarray((img.astype(np.uint32).sum(axis=2) * 80).astype(np.uint16)).save(buf, format="PNG")  # 16-bit greyscale
        with pytest.raises(rt.BackendError) as e:
            rt.decode_image(buf.getvalue())
        assert e.value.code == rt.RT_ERR_UNSUPPORTED
    good = bytearray(rt.encode_image(img, rt.IMAGE_PNG_RGBA8))
    good[60] ^= 0xFF  # corrupt IDAT: the chunk CRC no longer matches
    for bad in [bytes(good), bytes(good[:100]), b"JFIF not a png", b"P6\n2 2\n65535\n" + bytes(24)]:
        with pytest.raises(rt.BackendError):
            rt.decode_image(bad)

    earth = rng.integers(0, 256, size=(16, 32, 3), dtype=np.uint8)
    rt.set_earth_image(earth)
    try:
        flat = rt.FlatScene("book-two", 1)
        tex = flat.textures()
        images = tex[tex["kind"] == rt.TEX_IMAGE]
        assert len(images) == 1 and images[0]["width"] == 32 and images[0]["height"] == 16
        d = flat.desc()
        texels = np.ctypeslib.as_array(C.cast(d.texels, C.POINTER(C.c_float)), shape=(d.n_texels, 4))
        off = int(images[0]["texel_offset"])
        assert np.allclose(texels[off:off + 512, :3].reshape(16, 32, 3), earth / 255.0, atol=1e-7)
    finally:
        rt.set_earth_image(None)
    tex = rt.FlatScene("book-two", 1).textures()
    assert tex[tex["kind"] == rt.TEX_IMAGE][0]["width"] == 1024


def test_cpp_example_driver_builds_and_has_no_cpu_fallback():
    """examples/book_one.cpp (the C++ twin of rust/examples/book-one-gpu.rs) compiles against host/gpu.hpp and the C ABI; on a box
    without a CUDA device it must fail loudly (exit 1, the library's message), never render on the CPU."""
    pkg = os.path.join(ROOT, "ray-tracer_b200")
    subprocess.check_call(["make", "-s", "-C", pkg, "examples/book_one"])
    exe = os.path.join(pkg, "examples", "book_one")
    assert os.path.exists(exe)
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        has_gpu = False
    if not has_gpu:
        res = subprocess.run([exe, "8", "8", "1", "5", "1"], capture_output=True, text=True)
        assert res.returncode == 1 and "no CPU fallback" in res.stderr and not res.stdout
    res = subprocess.run([exe, "0", "8", "1", "5", "1"], capture_output=True, text=True)
    assert res.returncode == 2


def test_jpeg_decoder_matches_libjpeg_bit_for_bit(rt):
    """the ONE image format the reference opens is JPEG (examples/main.rs:266 `image::open("./earthmap.jpg")`).
    host/jpeg_decode.hpp (baseline + extended sequential Huffman; islow IDCT, fancy upsampling, fixed-point YCbCr) against PIL's
    libjpeg on 4:4:4 / 4:2:2 / 4:2:0 / grey streams, optimised Huffman tables, restart intervals, sizes that are not MCU multiples
    and a 1024x512 equirectangular-sized image: every byte equal.  Progressive streams are refused as unsupported, not mis-decoded."""
    import io
    PIL = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(1)

    def picture(h, w):
        yy, xx = np.mgrid[0:h, 0:w]
        img = np.stack([127 + 100 * np.sin(xx / 9.0) * np.cos(yy / 7.0), 40 + xx * (200.0 / w), 200 - yy * (150.0 / h)], axis=2) + rng.normal(0, 12, (h, w, 3))
        return np.clip(img, 0, 255).astype(np.uint8)

    cases = [((93, 141), dict(subsampling=0, quality=90)), ((93, 141), dict(subsampling=1, quality=85)), ((93, 141), dict(subsampling=2, quality=75)),
             ((64, 64), dict(subsampling=2, quality=30, optimize=True)), ((7, 5), dict(subsampling=2, quality=95)), ((1, 1), dict(quality=50)),
             ((512, 1024), dict(subsampling=2, quality=80)), ((50, 70), "grey"), ((37, 129), dict(subsampling=2, quality=60, restart_marker_blocks=3)),
             ((40, 40), dict(subsampling=0, quality=100))]
    for (h, w), kw in cases:
        img = picture(h, w)
        buf = io.BytesIO()
        if kw == "grey":
            PIL.fromarray(img[:, :, 0]).save(buf, format="JPEG", quality=80)
        else:
            try:
                PIL.fromarray(img).save(buf, format="JPEG", **kw)
            except TypeError:
                continue  # an older Pillow without restart_marker_blocks
        data = buf.getvalue()
        want = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
        got = rt.decode_image(data)
        assert got.shape == want.shape == (h, w, 3)
        assert np.array_equal(got, want), f"{h}x{w} {kw}: {np.abs(got.astype(int) - want.astype(int)).max()} off"
    buf = io.BytesIO()
    PIL.fromarray(picture(32, 32)).save(buf, format="JPEG", progressive=True)
    with pytest.raises(rt.BackendError) as e:
        rt.decode_image(buf.getvalue())
    assert e.value.code == rt.RT_ERR_UNSUPPORTED
    data = bytearray(data)
    with pytest.raises(rt.BackendError):
        rt.decode_image(bytes(data[:len(data) // 3]))  # truncated stream


def test_scene_plan_folds_axis_aligned_cubes_into_box_primitives(rt):
    """rt_scene_plan (host only): book-two's 400 ground blocks — six consecutive world-aligned rectangles each, every one a whole face —
    become 400 leaf items; the tree shrinks enough for extend's whole-tree-in-shared-memory form.  Rotated cubes (cornell-box is a
    small scene anyway) and scenes without rectangles are left alone."""
    plan = rt.FlatScene("book-two", 1).plan()
    assert plan["n_box_prims"] == 400 and plan["n_leaf_items"] == 3408 - 2400 + 400
    assert plan["whole_tree_threads"] == 1024 and plan["whole_tree_smem"] <= 224 * 1024
    o = rt.RtSceneOptions()
    o.flags = rt.SCENE_NO_BOX_PRIMS
    plain = rt.FlatScene("book-two", 1).plan(o)
    assert plain["n_box_prims"] == 0 and plain["n_leaf_items"] == 3408 and plain["whole_tree_threads"] == 0 and plain["n_nodes"] > 2 * plan["n_nodes"]
    one = rt.FlatScene("book-one", 1).plan()
    assert one["n_box_prims"] == 0 and one["whole_tree_threads"] == 512 and one["n_leaf_items"] == 485
    assert rt.FlatScene("cornell-box", 1).plan()["n_box_prims"] == 0  # a small scene: the frame-rect list handles its cubes


def _cube_rects(centre, half, first_order, material=0, rot=None):
    """The six face rectangles of a box, Cube::new style: [(R world<-local, centre, width, height)], normals outward."""
    import math

    def rot_axis(axis, deg):
        a = np.asarray(axis, dtype=np.float64)
        a /= np.linalg.norm(a)
        th = math.radians(deg)
        K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
        return np.eye(3) + math.sin(th) * K + (1 - math.cos(th)) * (K @ K)

    face = {"+x": rot_axis([0, 1, 0], 90), "-x": rot_axis([0, 1, 0], -90), "+y": rot_axis([1, 0, 0], -90), "-y": rot_axis([1, 0, 0], 90),
            "+z": np.eye(3), "-z": rot_axis([0, 1, 0], 180)}
    F = np.eye(3) if rot is None else rot_axis(*rot)
    out = []
    for name, sgn, ax in [("+x", 1, 0), ("-x", -1, 0), ("+y", 1, 1), ("-y", -1, 1), ("+z", 1, 2), ("-z", -1, 2)]:
        off = np.zeros(3)
        off[ax] = sgn * half[ax]
        ext = np.abs(face[name].T @ (2 * np.asarray(half, dtype=np.float64)))
        out.append((F @ face[name], np.asarray(centre, dtype=np.float64) + F @ off, ext[0], ext[1]))
    return out


def _rect_prims(rt, rects, orders=None):
    prims = np.zeros(len(rects), dtype=rt.PRIM_DTYPE)
    for i, (R, c, wd, ht) in enumerate(rects):
        prims[i]["kind"], prims[i]["material"], prims[i]["order"] = rt.PRIM_RECT, 0, i if orders is None else orders[i]
        prims[i]["param"][0], prims[i]["param"][1] = wd, ht
        prims[i]["world_from_local"] = np.concatenate([R, c[:, None]], axis=1).reshape(12).astype(np.float32)
        prims[i]["local_from_world"] = np.concatenate([R.T, (-R.T @ c)[:, None]], axis=1).reshape(12).astype(np.float32)
    mats = np.zeros(1, dtype=rt.MATERIAL_DTYPE)
    mats[0] = (rt.MAT_LAMBERTIAN, 0, 0, 0)
    texs = np.zeros(1, dtype=rt.TEXTURE_DTYPE)
    texs[0]["kind"], texs[0]["a"], texs[0]["b"] = rt.TEX_SOLID, -1, -1
    texs[0]["rgb"][:3] = (0.5, 0.5, 0.5)
    return prims, mats, texs


def test_box_primitive_detection_is_conservative(rt):
    """csrc/box_prims.h through rt_scene_plan: only six rectangles with CONSECUTIVE DFS ranks that are exactly the whole faces of one
    world-aligned box fold into a box primitive.  Rotated cubes, a cube with one face shrunk or missing, faces whose ranks are
    interleaved with another object's, and six coplanar-paired rectangles that do not close a box all stay rectangles."""
    filler = [_cube_rects([20.0 + 3 * k, 0, 0], [1, 1, 1], 0, rot=([1, 2, 3], 10.0 + k)) for k in range(6)]  # 36 rotated faces: never boxes, keep the scene > 32 prims

    def plan(cubes, orders=None):
        rects = [r for cube in cubes for r in cube] + [r for cube in filler for r in cube]
        if orders is not None:
            orders = list(orders) + list(range(len(orders), len(rects)))
        return rt.Scene.plan_tables(*_rect_prims(rt, rects, orders))

    n_fill = 36
    a = _cube_rects([0, 0, 0], [1, 2, 3], 0)
    b = _cube_rects([2, 0, 0], [1, 1, 1], 0)  # touches a's +x plane
    p = plan([a, b])
    assert p["n_box_prims"] == 2 and p["n_leaf_items"] == 2 + n_fill
    # rotated by 30 degrees: not world-aligned
    assert plan([_cube_rects([0, 0, 0], [1, 1, 1], 0, rot=([0, 1, 0], 30.0)), b])["n_box_prims"] == 1
    # rotated by exactly 90 degrees: still world-aligned (rows are signed unit vectors)
    assert plan([_cube_rects([0, 0, 0], [1, 2, 3], 0, rot=([0, 0, 1], 90.0))])["n_box_prims"] == 1
    # one face narrower than the box
    c = _cube_rects([0, 0, 0], [1, 1, 1], 0)
    R, ctr, wd, ht = c[2]
    c[2] = (R, ctr, wd * 0.9, ht)
    assert plan([c])["n_box_prims"] == 0
    # one face moved off its plane
    d = _cube_rects([0, 0, 0], [1, 1, 1], 0)
    R, ctr, wd, ht = d[5]
    d[5] = (R, ctr + np.array([0, 0, -0.01]), wd, ht)
    assert plan([d])["n_box_prims"] == 0
    # five faces only (an open box) followed by an unrelated rectangle
    e = _cube_rects([0, 0, 0], [1, 1, 1], 0)[:5] + [(np.eye(3), np.array([9.0, 9.0, 9.0]), 1.0, 1.0)]
    assert plan([e])["n_box_prims"] == 0
    # DFS ranks of two cubes interleaved: the faces are there but not as consecutive runs
    inter = [x for pair in zip(a, b) for x in pair]
    assert plan([inter])["n_box_prims"] == 0
    # same two cubes, table order shuffled but ranks consecutive per cube: detection goes by rank, not by table position
    rects = a + b
    perm = [7, 0, 9, 3, 11, 5, 1, 8, 2, 10, 4, 6]
    shuffled = [rects[j] for j in perm]
    assert plan([shuffled], orders=perm)["n_box_prims"] == 2
    # the option switches it off
    o = rt.RtSceneOptions()
    o.flags = rt.SCENE_NO_BOX_PRIMS
    rects = a + b + [r for cube in filler for r in cube]
    assert rt.Scene.plan_tables(*_rect_prims(rt, rects), options=o)["n_box_prims"] == 0
