"""The Rust side of the boundary (rust/src/gpu.rs, rust/build.rs) cannot be compiled in this image (no cargo / rustc), so it is
checked mechanically against what it binds: the `extern "C"` block against include/*.h (names, arity, argument types), the
#[repr(C)] structs against the C structs (field order, types, sizes), build.rs against ray-tracer_b200/Makefile (objects, arch,
link libraries), and — where /root/reference is present (this container, not the GPU box) — the trait impls against every
`impl Hit / Material / Texture for` of the reference and rust/reference.patch against its sources."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GPU_RS = os.path.join(ROOT, "rust", "src", "gpu.rs")
BUILD_RS = os.path.join(ROOT, "rust", "build.rs")
REF = "/root/reference"

C_TO_RUST = {"int": "c_int", "uint32_t": "u32", "int32_t": "i32", "uint64_t": "u64", "uint8_t": "u8", "float": "f32", "double": "f64", "void": "c_void",
             "char": "c_char"}


def _strip_c(src):
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def _c_type(t):
    """'const RtPrim*' -> '*const RtPrim'; 'RtScene**' -> '*mut *mut RtScene'; 'uint32_t' -> 'u32'."""
    t = t.strip()
    const = t.startswith("const ")
    t = t.replace("const ", "").strip()
    stars = t.count("*")
    base = t.replace("*", "").strip()
    base = C_TO_RUST.get(base, base)
    out = base
    for i in range(stars):
        out = ("*const " if (const and i == 0) else "*mut ") + out
    return out


def _c_functions():
    fns = {}
    for h in ("rt_b200.h", "rt_b200_host.h"):
        src = _strip_c(open(os.path.join(ROOT, "include", h)).read())
        for m in re.finditer(r"^\s*((?:const\s+)?\w+\s*\**)\s*(rth?_\w+)\s*\(([^)]*)\)\s*;", src, flags=re.M):
            ret, name, args = m.group(1), m.group(2), m.group(3)
            args = [a.strip() for a in args.split(",") if a.strip() and a.strip() != "void"]
            types = []
            for a in args:
                a = re.sub(r"\s+", " ", a)
                mm = re.match(r"(.*?)(\w+)$", a)  # strip the parameter name
                types.append(_c_type(mm.group(1)))
            fns[name] = (None if ret.strip() == "void" else _c_type(ret), types)
    return fns


def _rust_functions():
    src = re.sub(r"//[^\n]*", "", open(GPU_RS).read())
    block = re.search(r'extern "C" \{(.*?)\n    \}', src, flags=re.S).group(1)
    fns = {}
    for m in re.finditer(r"pub fn (\w+)\s*\((.*?)\)\s*(?:->\s*([^;]+))?;", block, flags=re.S):
        name, args, ret = m.group(1), m.group(2), m.group(3)
        types = [re.sub(r"\s+", " ", a.split(":", 1)[1].strip()) for a in re.split(r",\s*(?=\w+\s*:)", args.strip()) if a.strip()]
        fns[name] = (ret.strip() if ret else None, types)
    return fns


def test_extern_block_matches_the_headers():
    c, r = _c_functions(), _rust_functions()
    core = {n for n in c if n.startswith("rt_")}
    assert core <= set(r), f"rt_b200.h entry points not bound by gpu.rs: {sorted(core - set(r))}"
    assert set(r) <= set(c), f"gpu.rs binds symbols the headers do not declare: {sorted(set(r) - set(c))}"
    assert len(r) >= 27
    for name, (ret, types) in r.items():
        cret, ctypes_ = c[name]
        assert len(types) == len(ctypes_), f"{name}: arity {len(types)} in gpu.rs, {len(ctypes_)} in the header"
        assert ret == cret, f"{name}: returns {ret} in gpu.rs, {cret} in the header"
        for i, (a, b) in enumerate(zip(types, ctypes_)):
            assert a == b, f"{name}: argument {i} is `{a}` in gpu.rs, `{b}` in the header"


def _c_structs():
    src = _strip_c(open(os.path.join(ROOT, "include", "rt_b200.h")).read())
    out = {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", src, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = re.sub(r"\s+", " ", decl).strip()
            if not decl:
                continue
            mm = re.match(r"((?:const )?\w+\s*\**)\s*(.*)$", decl)
            base = mm.group(1)
            for item in mm.group(2).split(","):
                item = item.strip()
                arr = re.match(r"(\w+)\[(\d+)\]$", item)
                if arr:
                    fields.append((arr.group(1), f"[{_c_type(base)}; {arr.group(2)}]"))
                else:
                    stars = item.count("*")
                    fields.append((item.replace("*", "").strip(), _c_type(base + "*" * stars)))
        out[m.group(1)] = fields
    return out


def _rust_structs():
    src = re.sub(r"//[^\n]*", "", open(GPU_RS).read())
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^)]*\)\]\s*)?pub struct (\w+) \{(.*?)\n    \}", src, flags=re.S):
        fields = [(f.group(1), re.sub(r"\s+", " ", f.group(2).strip())) for f in re.finditer(r"pub (\w+): ([^,\n]+),", m.group(2))]
        out[m.group(1)] = fields
    return out


def test_repr_c_structs_match_the_header(tmp_path):
    c, r = _c_structs(), _rust_structs()
    names = ["RtPrim", "RtMaterial", "RtTexture", "RtSceneDesc", "RtCamera", "RtRenderParams", "RtStats", "RtRayIn", "RtHitOut", "RtContextOptions",
             "RtSceneOptions"]
    for n in names:
        assert n in c and n in r, n
        assert r[n] == c[n], f"{n}: gpu.rs {r[n]} vs header {c[n]}"
    # sizes the comments in gpu.rs promise, from the C compiler
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include "rt_b200.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(RtPrim), sizeof(RtMaterial), '
                    "sizeof(RtTexture), sizeof(RtCamera), sizeof(RtRayIn), sizeof(RtHitOut), sizeof(RtContextOptions), sizeof(RtSceneOptions));return 0;}")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    assert subprocess.check_output([str(exe)]).split() == [b"128", b"16", b"48", b"64", b"24", b"48", b"32", b"32"]
    src = open(GPU_RS).read()
    assert re.search(r"RT_ABI_VERSION: u32 = (\d+)", src).group(1) == re.search(r"#define RT_ABI_VERSION (\d+)u", open(os.path.join(ROOT, "include", "rt_b200.h")).read()).group(1)


def test_build_rs_lists_the_makefile_objects_and_link_line():
    mk = open(os.path.join(ROOT, "ray-tracer_b200", "Makefile")).read()
    objs = re.search(r"^OBJ := (.*)$", mk, flags=re.M).group(1).split()
    rs = open(BUILD_RS).read()
    units = re.findall(r'\("([\w/.]+)", "(\w+\.o)", (true|false)\)', rs)
    assert sorted(o for _, o, _ in units) == sorted(os.path.basename(o) for o in objs)
    for src, obj, is_cuda in units:
        assert os.path.exists(os.path.join(ROOT, "ray-tracer_b200", src)), src
        assert (is_cuda == "true") == src.endswith(".cu")
        rule = re.search(rf"^build/{re.escape(obj)}: (\S+)", mk, flags=re.M).group(1)
        assert rule == src
    assert "arch=compute_100a,code=sm_100a" in rs and "arch=compute_100a,code=sm_100a" in mk
    link = re.search(r"-shared -o \$@ \$\(OBJ\) (.*)$", mk, flags=re.M).group(1)
    assert "-cudart static" in link and "cudart_static" in rs
    for lib in re.findall(r"-l(\w+)", link):
        assert f'"{lib}"' in rs, f"build.rs does not link {lib}"
    for h in re.findall(r'"((?:csrc|host|examples|\.\./include)/[\w./]+)"', rs):
        assert os.path.exists(os.path.join(ROOT, "ray-tracer_b200", h)), h
    hdr = re.search(r"^HDR := (.*?)\n\n", mk, flags=re.S | re.M).group(1).replace("\\\n", " ").split()
    listed = set(re.findall(r'"((?:csrc|host|examples|\.\./include)/[\w./]+\.(?:h|hpp|cuh))"', rs))
    assert listed == set(hdr), (sorted(listed ^ set(hdr)))


@pytest.mark.skipif(not os.path.isdir(REF), reason="the reference sources exist in the build container only")
def test_every_reference_implementor_has_its_hook_and_the_patch_applies(tmp_path):
    rs = re.sub(r"//[^\n]*", "", open(GPU_RS).read())

    def impls(trait):
        found = []
        for f in os.listdir(os.path.join(REF, "src")):
            for m in re.finditer(rf"^impl(?:<[^>]*>)?\s+{trait}\s+for\s+(.+?)\s*(?:\{{|$)", open(os.path.join(REF, "src", f)).read(), flags=re.M):
                found.append(re.sub(r"\s+", " ", m.group(1)))
        return found

    def hooked(hook, ty):
        head = re.sub(r"<.*", "", ty)  # Sprite<T, U> -> Sprite
        if ty.startswith("Vec<"):
            return re.search(rf"impl {hook} for {re.escape(ty)}", rs) is not None
        return re.search(rf"impl(?:<[^>]*>)?\s+{hook}\s+for\s+{re.escape(head)}\b", rs) is not None

    for trait, hook in (("Hit", "Flatten"), ("Material", "Describe"), ("Texture", "DescribeTexture")):
        types = impls(trait)
        assert len(types) >= 3
        for ty in types:
            assert hooked(hook, ty), f"`impl {trait} for {ty}` in the reference has no `impl {hook} for ..` in gpu.rs"
    # every reference getter gpu.rs calls exists upstream or is added by the patch
    dst = tmp_path / "ref"
    shutil.copytree(REF, dst)
    patch = "\n".join(ln for ln in open(os.path.join(ROOT, "rust", "reference.patch")).read().splitlines() if not ln.startswith("#")) + "\n"
    subprocess.run(["patch", "-p1", "--forward"], input=patch.encode(), cwd=dst, check=True, stdout=subprocess.DEVNULL)
    patched = {f: open(os.path.join(dst, "src", f)).read() for f in os.listdir(os.path.join(dst, "src"))}
    everything = "\n".join(patched.values())
    for getter in ["radius", "width", "height", "geometry", "transform", "material", "boundary", "density", "left", "right", "albedo", "fuzziness",
                   "refractive", "emission", "black", "white", "origin", "inversed", "multiplied", "eye", "center", "up", "fov", "aspect", "focusDistance",
                   "lensRadius", "direction"]:
        assert re.search(rf"pub fn {getter}\b", everything), getter
    assert "crate::gpu::Flatten" in patched["ray.rs"] and "pub mod gpu;" in patched["lib.rs"]
    assert "crate::gpu::Describe" in patched["material.rs"] and "crate::gpu::DescribeTexture" in patched["material.rs"]


def test_rust_constants_match_the_header_enumerators():
    """Every `pub const RT_*` of gpu.rs carries the value include/rt_b200.h gives that enumerator / macro (flags, kinds, error codes,
    the rt_render_noalloc limits)."""
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "rt_b200.h")).read()
    rs = open(os.path.join(root, "rust", "src", "gpu.rs")).read()

    def value(text):
        text = text.strip().rstrip("uU").replace("_", "")
        m = re.fullmatch(r"\(?\s*(\d+)u?\s*<<\s*(\d+)\s*\)?", text)
        if m:
            return int(m.group(1)) << int(m.group(2))
        return int(text, 0)

    c_vals = {}
    for name, v in re.findall(r"\b(RT_[A-Z0-9_]+)\s*=\s*(-?(?:0x[0-9A-Fa-f]+|\d+))u?\s*[,/}\n]", hdr):
        c_vals[name] = value(v)
    for name, v in re.findall(r"#define[ \t]+(RT_[A-Z0-9_]+)[ \t]+(\(?[0-9xA-Fa-fuU< \t]+\)?)[ \t]*(?:/\*.*)?$", hdr, flags=re.M):
        c_vals[name] = value(v)
    rust = re.findall(r"pub const (RT_[A-Z0-9_]+):\s*\w+\s*=\s*([^;]+);", rs)
    assert len(rust) >= 40
    checked = 0
    for name, v in rust:
        assert name in c_vals, f"{name} is not declared in include/rt_b200.h"
        assert value(v) == c_vals[name], (name, v, c_vals[name])
        checked += 1
    assert checked == len(rust)
