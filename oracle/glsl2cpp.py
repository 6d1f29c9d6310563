#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — build-time GLSL -> C++20(+glm) text rewrite for the oracle.

Reads one of the reference's ray-tracing shader sources *in place* from $BLOK_REF
(assets/shaders/{intersect.rint, raygen.rgen, hit.rchit}) and writes a C++ fragment into
oracle/_ref/gen/ (git-ignored; never committed, never shipped in the product). The fragment
is #included inside a namespace by oracle/ref_harness.cpp, so the arithmetic executed on the
CPU *is the reference's source*, not a paraphrase.

The rewrite is purely mechanical:
  * drop #version / #extension lines
  * storage buffers  -> `static const T* name;`
  * uniform block    -> `struct FrameUBO_t {...}; static FrameUBO_t frame;`
  * storage images   -> `static image2D name;`
  * rayPayload[In]EXT / hitAttributeEXT globals -> `static thread_local T name;`
  * `inout T x` -> `T& x`; `void main()` -> `void shader_main()`
  * float literals get an `f` suffix (GLSL literals are fp32)
  * multi-component swizzles `.xyz` -> `.xyz()` (glm with GLM_FORCE_SWIZZLE)
  * GLSL array constructor `vec3[]( ... )` -> `{ ... }`
Additive instrumentation (does not change any arithmetic): capture the node index passed with
reportIntersectionEXT, count node fetches, count MAX_ITER cap hits.
"""
import re
import sys


def _array_ctor(src: str) -> str:
    # `T[]( a, b, c )` -> `{ a, b, c }` (balanced parentheses)
    out = []
    i = 0
    pat = re.compile(r"\b\w+\[\d*\]\(")
    while True:
        m = pat.search(src, i)
        if not m:
            out.append(src[i:])
            break
        out.append(src[i:m.start()])
        depth = 1
        j = m.end()
        while depth:
            c = src[j]
            depth += (c == "(") - (c == ")")
            j += 1
        out.append("{" + src[m.end():j - 1] + "}")
        i = j
    return "".join(out)


def rewrite_comp(src: str) -> str:
    """Compute shaders (temporal_reproject / variance / atrous / taa / tonemap / sharpen .comp): same mechanical rules,
    plus storage images with format qualifiers, combined image samplers, push constants and `out` parameters.
    Appends `bind_image(binding, image)` so the harness can attach buffers by the shader's own binding numbers."""
    src = re.sub(r"^\s*#(version|extension).*$", "", src, flags=re.M)
    src = re.sub(r"layout\(local_size_x[^)]*\)\s*in\s*;", "", src)
    binds = []

    def img(m):
        binds.append((int(m.group(1)), m.group(3), m.group(2)))
        return f"static {m.group(2)} {m.group(3)};"
    src = re.sub(r"layout\(\s*binding\s*=\s*(\d+)[^)]*\)\s*uniform\s+(?:readonly\s+|writeonly\s+)?(image2D|sampler2D)\s+(\w+)\s*;", img, src)
    src = re.sub(r"layout\([^)]*\)\s*uniform\s+FrameUBO\s*\{", "struct FrameUBO_t {", src)
    src = re.sub(r"\}\s*(frame|ubo)\s*;", r"};\nstatic FrameUBO_t \1;", src)
    src = re.sub(r"layout\(push_constant\)\s*uniform\s+PushConstants\s*\{", "struct PushConstants_t {", src)
    src = re.sub(r"\}\s*pc\s*;", "};\nstatic PushConstants_t pc;", src)
    src = re.sub(r"\binout\s+(\w+)\s+(\w+)", r"\1& \2", src)
    src = re.sub(r"\bout\s+(\w+)\s+(\w+)", r"\1& \2", src)
    src = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", src)
    src = re.sub(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])", r"\1f", src)
    src = re.sub(r"\.(xyz|xy|rgb|rg)\b(?!\s*\()", r".\1()", src)
    src = _array_ctor(src)
    if not binds or "shader_main" not in src:
        raise SystemExit("glsl2cpp: compute shader anchors not found")
    src += "\nstatic bool bind_image(int binding, const image2D& im) {\n    switch (binding) {\n"
    for b, name, ty in sorted(binds):
        src += f"        case {b}: {name}{'.img' if ty == 'sampler2D' else ''} = im; return true;\n"
    src += "    }\n    return false;\n}\n"
    return src


def rewrite(src: str, kind: str) -> str:
    if kind == "comp":
        return rewrite_comp(src)
    src = re.sub(r"^\s*#(version|extension).*$", "", src, flags=re.M)
    # storage buffers
    src = re.sub(r"layout\([^)]*\)\s*readonly\s+buffer\s+\w+\s*\{\s*(\w+)\s+(\w+)\[\];\s*\};",
                 r"static const \1* \2;", src)
    # the frame uniform block
    src = re.sub(r"layout\([^)]*\)\s*uniform\s+FrameUBO\s*\{", "struct FrameUBO_t {", src)
    src = re.sub(r"\}\s*frame\s*;", "};\nstatic FrameUBO_t frame;", src)
    # acceleration structure + storage images
    src = re.sub(r"layout\([^)]*\)\s*uniform\s+accelerationStructureEXT\s+(\w+);",
                 r"static accelerationStructureEXT \1;", src)
    src = re.sub(r"layout\([^)]*\)\s*uniform\s+image2D\s+(\w+);", r"static image2D \1;", src)
    # payloads / hit attributes (per-invocation state -> thread_local)
    src = re.sub(r"layout\(location\s*=\s*\d+\)\s*rayPayload(?:In)?EXT\s+(\w+)\s+(\w+);",
                 r"static thread_local \1 \2;", src)
    src = re.sub(r"hitAttributeEXT\s+(\w+)\s+(\w+);", r"static thread_local \1 \2;", src)
    # qualifiers
    src = re.sub(r"\binout\s+(\w+)\s+(\w+)", r"\1& \2", src)
    src = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", src)
    # float literals -> fp32
    src = re.sub(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])", r"\1f", src)
    # swizzles
    src = re.sub(r"\.(xyz|xy|rgb|rg)\b(?!\s*\()", r".\1()", src)
    src = _array_ctor(src)
    # GLSL evaluates constructor arguments left to right; C++ leaves call-argument order unspecified.
    # The one constructor whose arguments have side effects (two RNG draws) becomes a braced init, which
    # C++ does sequence left to right.
    src = src.replace("return vec2(randomFloat(state), randomFloat(state));",
                      "return vec2{randomFloat(state), randomFloat(state)};")
    if kind == "rint":
        n0 = src.count("reportIntersectionEXT(")
        src = src.replace("reportIntersectionEXT(", "(ref_cap_node = item.nodeIndex), reportIntersectionEXT(")
        src, n1 = re.subn(r"(SvoNode\s+node\s*=\s*nodes\[item\.nodeIndex\];)", r"\1 ref_cnt_fetch++;", src)
        src, n2 = re.subn(r"iter\+\+\s*<\s*MAX_ITER", "(iter++ < MAX_ITER || (ref_cnt_cap++, false))", src)
        src, n3 = re.subn(r"(if\s*\(tMin\s*>\s*tMax\)\s*return;)", r"\1 ref_cnt_aabb++;", src)
        if not (n0 == 1 and n1 == 1 and n2 == 1 and n3 == 1):
            raise SystemExit(f"glsl2cpp: instrumentation anchors not found ({n0},{n1},{n2},{n3})")
    return src


if __name__ == "__main__":
    inp, outp, kind = sys.argv[1:4]
    with open(inp) as f:
        text = f.read()
    with open(outp, "w") as f:
        f.write(f"// GENERATED at build time from {inp} — do not commit, do not ship.\n")
        f.write(rewrite(text, kind))
