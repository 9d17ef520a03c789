#!/usr/bin/env python
"""Generates rust/orbx-sys/src/lib.rs from include/orbx.h, so that the raw FFI crate is a 1:1 mirror of the C ABI by
construction (tests/test_abi_cpu.py diffs the symbol sets of the header, the ctypes table, the built library and this
file).  Usage: python scripts/gen_rust_sys.py [--check]"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "orbx.h")
OUT = os.path.join(ROOT, "rust", "orbx-sys", "src", "lib.rs")

SCALARS = {"int": "c_int", "float": "f32", "double": "f64", "size_t": "usize", "int32_t": "i32", "int64_t": "i64",
           "uint8_t": "u8", "uint32_t": "u32", "uint64_t": "u64", "char": "c_char", "void": "c_void"}
KEYWORDS = {"in", "type", "match", "ref", "move", "box", "fn", "use", "mod", "self", "loop"}


def strip_comments(text):
    return re.sub(r"/\*.*?\*/", " ", text, flags=re.S)


def rust_type(ctype):
    t = " ".join(ctype.replace("*", " * ").split())
    parts = t.split(" ")
    const = parts[0] == "const"
    if const:
        parts = parts[1:]
    base, stars = parts[0], parts.count("*")
    if base in ("unsigned",):
        raise ValueError("unsupported C type " + ctype)
    r = SCALARS.get(base, base)  # struct / opaque names are emitted verbatim
    for i in range(stars):
        # only the innermost level of a C declaration in this header is ever const
        r = ("*const " if (const and i == 0) else "*mut ") + r
    return r


def parse(text):
    text = strip_comments(text)
    defines = re.findall(r"#define\s+(ORBX_\w+)\s+\(?(-?\d+)\)?", text)
    enums = []
    for body in re.findall(r"enum\s*\{([^}]*)\}", text):
        for name, val in re.findall(r"(ORBX_\w+)\s*=\s*(-?\d+)", body):
            enums.append((name, val))
    structs = []
    for body, name in re.findall(r"typedef\s+struct\s*\{([^}]*)\}\s*(\w+)\s*;", text, flags=re.S):
        fields = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            ty, names = decl.split(" ", 1)
            for nm in names.split(","):
                fields.append((nm.strip(), SCALARS[ty]))
        structs.append((name, fields))
    opaque = re.findall(r"typedef\s+struct\s+(\w+)\s+\1\s*;", text)
    funcs = []
    for ret, name, args in re.findall(r"\n\s*((?:const\s+)?\w+\s*\**)\s*(orbx_\w+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        params = []
        args = " ".join(args.split())
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                m = re.match(r"(.*?)(\w+)$", a)
                ty, nm = m.group(1).strip(), m.group(2)
                if nm in KEYWORDS:
                    nm += "_"
                params.append((nm, rust_type(ty)))
        r = " ".join(ret.split())
        funcs.append((name, params, None if r == "void" else rust_type(r)))
    return defines, enums, structs, opaque, funcs


def generate():
    defines, enums, structs, opaque, funcs = parse(open(HEADER).read())
    out = ["//! Raw FFI to liborbx: GENERATED from include/orbx.h by scripts/gen_rust_sys.py - do not edit.",
           "//! Every declaration of the header appears here once; the safe wrappers live in the `orbx` crate.",
           "#![allow(non_camel_case_types)]", "use std::os::raw::{c_char, c_int, c_void};", ""]
    for name, fields in structs:
        derive = "Clone, Copy, Debug" + (", Default" if name != "orbx_params" else "")
        out.append("#[repr(C)]\n#[derive(%s)]" % derive)
        out.append("pub struct %s {" % name)
        for nm, ty in fields:
            out.append("    pub %s: %s," % (nm, ty))
        out.append("}")
    for name in opaque:
        out.append("pub enum %s {}" % name)
    out.append("")
    for name, val in defines + enums:
        out.append("pub const %s: c_int = %s;" % (name, val))
    out.append("")
    out.append('extern "C" {')
    for name, params, ret in funcs:
        sig = ", ".join("%s: %s" % p for p in params)
        out.append("    pub fn %s(%s)%s;" % (name, sig, " -> " + ret if ret else ""))
    out.append("}")
    return "\n".join(out) + "\n", [f[0] for f in funcs]


if __name__ == "__main__":
    text, names = generate()
    if "--check" in sys.argv:
        cur = open(OUT).read() if os.path.exists(OUT) else ""
        sys.exit(0 if cur == text else 1)
    open(OUT, "w").write(text)
    print("wrote %s: %d functions" % (OUT, len(names)))
