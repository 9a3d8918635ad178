#!/usr/bin/env python3
"""Prepare the reference's GLSL sources for compilation as C++ (oracle/_ref build step).

Reads the shader sources where they lie under <reference>/shaders and writes mechanically transformed copies
into oracle/_ref/gen/ (git-ignored build output; reference sources are never committed to this repo).

Transformations (purely lexical, no arithmetic is touched):
  * drop `#version`, `#extension` and `layout(local_size...) in;` lines;
  * give every floating-point literal an `f` suffix: GLSL literals without suffix are 32-bit floats, C++ ones
    are doubles, and per-operation FP32 rounding is what the parity contract is about;
  * `inout T x` / `out T x` parameters become `T& x`; the `in` qualifier is dropped;
  * swizzles used by the shaders (`.xy`, `.rgb`) become accessor calls;
  * GLSL evaluates function arguments left to right (GLSL 4.50 section 6.1.1); g++ evaluates them right to left.
    The one call whose two arguments both have side effects, `randomDir(randf_inc(seed), randf_inc(seed))`
    (light_gen.comp:20,72; path_compute_color.comp:97), is therefore sequenced explicitly with a statement expression.
common_bindings.h (descriptor declarations only) and PNanoVDB_GLSL.h are NOT copied: the shim provides
`common_bindings.h` (plain members) and routes pnanovdb_* calls to the vendored upstream PNanoVDB.h in C mode.
"""
import os
import re
import sys

FILES = [
    "common_structures.h", "common_functions.h", "common_nanovdb.h", "common_includes.h",
    "light_gen.comp", "point_light_gen.comp", "ray_light_gen.comp", "sphere_light_gen.comp", "beam_light_gen.comp",
    "path_light_gen.comp",
    "point_compute_color.comp", "sphere_compute_color.comp", "ray_compute_color.comp", "beam_compute_color.comp",
    "path_compute_color.comp",
]

FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")


def transform(text: str) -> str:
    out = []
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("#version") or s.startswith("#extension"):
            continue
        if re.match(r"layout\s*\(\s*local_size", s):
            continue
        code, sep, comment = line.partition("//")
        code = FLOAT_LIT.sub(lambda m: m.group(1) + "f", code)
        code = re.sub(r"\binout\s+(\w+)\s+", r"\1& ", code)
        code = re.sub(r"\bout\s+(\w+)\s+", r"\1& ", code)
        code = re.sub(r"(?<=[(,])\s*in\s+(\w+)\s+", r" \1 ", code)
        code = code.replace("randomDir(randf_inc(seed), randf_inc(seed))",
                            "({ float lr_a = randf_inc(seed); float lr_b = randf_inc(seed); randomDir(lr_a, lr_b); })")
        code = re.sub(r"\.xy\b", ".xy()", code)
        code = re.sub(r"\.rgb\b", ".rgb()", code)
        out.append(code + sep + comment)
    return "\n".join(out) + "\n"


def main():
    ref_shaders, out_dir = sys.argv[1], sys.argv[2]
    os.makedirs(out_dir, exist_ok=True)
    for name in FILES:
        with open(os.path.join(ref_shaders, name), "r", encoding="utf-8", errors="replace") as f:
            text = f.read()
        with open(os.path.join(out_dir, name), "w") as f:
            f.write(transform(text))
    # PNanoVDB_GLSL.h is replaced by the shim's wrappers (already included at namespace scope)
    with open(os.path.join(out_dir, "PNanoVDB_GLSL.h"), "w") as f:
        f.write("/* intentionally empty: see oracle/ref/glsl_shim.h */\n")


if __name__ == "__main__":
    main()
