"""The Rust shim crate (raytracer-cuda/) cannot be compiled here (no rustc/cargo in the image). What can be checked
without a compiler is checked: src/ffi.rs against include/rtc.h — every exported function with its argument count,
every struct with its field names, order and size — that the crate's files exist, and that
patches/raytracer-accessors.diff applies cleanly to the reference tree (when /root/reference is present)."""
import ctypes as C
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CRATE = os.path.join(ROOT, "raytracer-cuda")


def _strip_c(text):
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"//[^\n]*", "", text)


def header_functions():
    """name -> number of parameters, from the RTC_API declarations of include/rtc.h"""
    text = _strip_c(open(os.path.join(ROOT, "include", "rtc.h")).read())
    out = {}
    for m in re.finditer(r"RTC_API\s+[^;(]*?\b(rtc_[a-z0-9_]+)\s*\(([^;]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len([a for a in args.split(",") if a.strip()])
    return out


def header_structs():
    """name -> [(field, ctype, array_len)] in declaration order"""
    text = _strip_c(open(os.path.join(ROOT, "include", "rtc.h")).read())
    out = {}
    for m in re.finditer(r"typedef\s+struct\s+(rtc_[a-z0-9_]+)\s*\{(.*?)\}\s*\1\s*;", text, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            ty, names = re.match(r"((?:const\s+)?[A-Za-z0-9_]+\s*\*?)\s*(.*)", decl, flags=re.S).groups()
            ty = ty.replace(" ", "")
            for nm in names.split(","):
                nm = nm.strip()
                ptr = nm.startswith("*") or ty.endswith("*")
                nm = nm.lstrip("*").strip()
                arr = re.match(r"(\w+)\[(\d+)\]", nm)
                if arr:
                    fields.append((arr.group(1), ty, int(arr.group(2)), ptr))
                else:
                    fields.append((nm, ty, 0, ptr))
        out[m.group(1)] = fields
    return out


def rust_source():
    text = open(os.path.join(CRATE, "src", "ffi.rs")).read()
    return re.sub(r"//[^\n]*", "", text)


def rust_functions():
    text = rust_source()
    block = re.search(r'extern\s+"C"\s*\{(.*?)\n\}', text, flags=re.S).group(1)
    out = {}
    for m in re.finditer(r"pub\s+fn\s+(rtc_[a-z0-9_]+)\s*\((.*?)\)\s*(?:->\s*[^;]+)?;", block, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if not args else len([a for a in args.split(",") if a.strip()])
    return out


def rust_structs():
    text = rust_source()
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\][^{]*?pub\s+struct\s+(rtc_[a-z0-9_]+)\s*\{(.*?)\n\}", text, flags=re.S):
        fields = []
        for f in re.finditer(r"pub\s+(\w+)\s*:\s*([^,\n]+),", m.group(2)):
            fields.append((f.group(1), f.group(2).strip()))
        out[m.group(1)] = fields
    return out


RUST_SIZE = {"i32": 4, "u32": 4, "f64": 8, "u64": 8, "u8": 1}
C_TO_RUST = {"int32_t": "i32", "uint32_t": "u32", "double": "f64", "uint64_t": "u64"}


def rust_layout(fields):
    """size of a #[repr(C)] struct with natural alignment"""
    off, align = 0, 1
    for _, ty in fields:
        arr = re.match(r"\[(\w+);\s*(\d+)\]", ty)
        if ty.startswith("*"):
            sz, al, n = 8, 8, 1
        elif arr:
            sz = al = RUST_SIZE[arr.group(1)]
            n = int(arr.group(2))
        else:
            sz = al = RUST_SIZE[ty]
            n = 1
        off = (off + al - 1) // al * al
        off += sz * n
        align = max(align, al)
    return (off + align - 1) // align * align


def test_crate_files_exist():
    for f in ["Cargo.toml", "build.rs", "README.md", "src/lib.rs", "src/ffi.rs", "src/flatten.rs", "patches/raytracer-accessors.diff"]:
        assert os.path.isfile(os.path.join(CRATE, f)), f
    assert "Not compiled here" in open(os.path.join(CRATE, "README.md")).read()
    build = open(os.path.join(CRATE, "build.rs")).read()
    assert "arch=compute_100a,code=sm_100a" in build and "--fmad=false" in build


def test_ffi_binds_every_export(pkg):
    hdr, rs = header_functions(), rust_functions()
    assert len(hdr) >= 23
    assert sorted(hdr) == sorted(rs), (sorted(set(hdr) - set(rs)), sorted(set(rs) - set(hdr)))
    for name, n in hdr.items():
        assert rs[name] == n, (name, n, rs[name])
        assert hasattr(pkg.cuda_lib(), name)


def test_ffi_struct_fields_and_sizes(pkg):
    hdr, rs = header_structs(), rust_structs()
    want = ["rtc_node", "rtc_material", "rtc_pattern", "rtc_light", "rtc_scene_desc", "rtc_camera", "rtc_opts", "rtc_stats"]
    for name in want:
        assert name in hdr and name in rs, name
        hf, rf = hdr[name], rs[name]
        assert [f[0] for f in hf] == [f[0] for f in rf], name
        for (hn, hty, harr, hptr), (rn, rty) in zip(hf, rf):
            if hptr:
                assert rty.startswith("*const"), (name, hn, rty)
            elif harr:
                assert rty == "[%s; %d]" % (C_TO_RUST[hty], harr), (name, hn, rty)
            else:
                assert rty == C_TO_RUST[hty], (name, hn, rty)
    sizes = {"rtc_node": pkg.RtcNode, "rtc_camera": pkg.RtcCamera, "rtc_opts": pkg.RtcOpts, "rtc_stats": pkg.RtcStats,
             "rtc_scene_desc": pkg.RtcSceneDesc}
    for name, ct in sizes.items():
        assert rust_layout(rs[name]) == C.sizeof(ct), name
    assert rust_layout(rs["rtc_material"]) == 88 and rust_layout(rs["rtc_pattern"]) == 152 and rust_layout(rs["rtc_light"]) == 48


def test_ffi_constants_match_header():
    text = open(os.path.join(ROOT, "include", "rtc.h")).read()
    rs = rust_source()
    for name in re.findall(r"\b(RTC_(?:ERR_[A-Z]+|OK|SPHERE|PLANE|CUBE|CYLINDER|CONE|TRIANGLE|SMOOTH_TRIANGLE|GROUP|CSG|PAT_[A-Z]+|CSG_[A-Z]+))\s*=\s*(-?\d+)", text):
        m = re.search(r"pub const %s: \w+ = (-?\d+);" % name[0], rs)
        assert m and int(m.group(1)) == int(name[1]), name
    for name, val in re.findall(r"#define\s+(RTC_(?:FLAG|OPT|STAT)_[A-Z_]+|RTC_ABI_VERSION|RTC_TRACE_STATIC_MAX)\s+(\d+)u?", text):
        m = re.search(r"pub const %s: \w+ = (\d+);" % name, rs)
        assert m and int(m.group(1)) == int(val), name


@pytest.mark.skipif(not os.path.isdir("/root/reference/raytracer/src") or shutil.which("patch") is None,
                    reason="needs the reference tree and patch(1)")
def test_accessor_patch_applies_to_the_reference(tmp_path):
    dst = tmp_path / "ref"
    shutil.copytree("/root/reference", dst, ignore=shutil.ignore_patterns(".git", "target"))
    subprocess.run(["chmod", "-R", "u+w", str(dst)], check=True)
    p = subprocess.run(["patch", "-p1", "--dry-run", "-i", os.path.join(CRATE, "patches", "raytracer-accessors.diff")],
                       cwd=dst, capture_output=True, text=True)
    assert p.returncode == 0, p.stdout + p.stderr
    # every accessor src/flatten.rs and src/lib.rs call is one the patch adds or the reference already has
    patch = open(os.path.join(CRATE, "patches", "raytracer-accessors.diff")).read()
    for fn in ["objects", "lights", "hsize", "vsize", "pixel_size", "half_width", "half_height", "transform_inverse",
               "get_num_threads", "get_aa_samples", "pattern", "kind_and_colors", "minimum", "maximum", "closed", "operation"]:
        assert re.search(r"^\+\s+pub fn %s\(" % fn, patch, flags=re.M), fn
