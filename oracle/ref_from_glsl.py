#!/usr/bin/env python3
"""ORACLE — TEST INFRASTRUCTURE ONLY.

Source-to-source recipe: reads the REFERENCE'S OWN shader text (shaders/raytrace.comp) and rewrites it,
mechanically and line by line, into a C++ fragment that compiles against the reference's vendored GLM 0.9.9.7
(third_party/glm, GLM_FORCE_SWIZZLE). The fragment is written to oracle/_ref/raytrace_comp.inc (git-ignored: it
is the reference's text, never committed) and is #included by oracle/ref_shim.cpp, which supplies the GLSL
built-ins the fragment calls and a plain-C API. The result, oracle/_ref/libraytrace_comp.so, is "the reference
shader run on the CPU"; tests/test_ref_shader.py asserts that the hand-written oracle (oracle_shader.cpp) equals
it bit for bit.

Every rewrite is a syntactic rule listed in RULES below; none of them changes an arithmetic statement:
  R1  drop `#version` and the `layout(local_size...) in;` line
  R2  `layout(... ) readonly buffer B { T name[]; };`        -> `static const T* name;` (set by the shim)
  R3  `layout(push_constant) uniform constants {...} PushConstants;` -> `static struct constants {...} PushConstants;`
  R4  `layout(binding = 0, rgba8) uniform image2D outImage;` -> `static image2D outImage;` (shim type)
      texture / sampler array declarations (never referenced by the HEAD shader) are dropped
  R5  GLSL array declarators `T[N] name` -> `T name[N]`
  R6  `inout T name[N]` -> `T name[N]` (a C++ array parameter is already by reference), `inout T name` -> `T& name`
  R7  floating literals without suffix get `f` (every GLSL float literal is 32-bit)
  R8  scalar / vector / struct locals and struct members declared without an initialiser are value-initialised
      (`T x;` -> `T x{};`): GLSL leaves them undefined; SURVEY A-Q9 pins the one that matters (HitInfo.materialIndex on a
      miss) to 0. Local ARRAYS (`uint stack[64]`, the throw-away `float[2]` counters) stay uninitialised as in GLSL: the
      shader writes stack[0] before it reads, and the throw-away counters are never consumed
  R9  `void main()` -> `void glsl_main()`
  R10 `const float` globals get `static`
Implicit int -> float conversions of GLSL (`max(0, x)`, `vec3 * int`, `pow(x, 5)`, `smoothstep(0, 0.4, t)`) need no
rewrite: the shim declares the built-ins with float parameters inside the same namespace, and C++ converts.
"""
import os
import re
import sys

REF_SHADER = "/root/reference/shaders/raytrace.comp"

FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])")
STRUCT_TYPES = set()
BASIC_TYPES = {"float", "int", "uint", "bool", "vec2", "vec3", "vec4", "uvec3", "ivec2", "mat3", "mat4"}


def strip_comment(line):
    i = line.find("//")
    return (line, "") if i < 0 else (line[:i], line[i:])


def translate(src):
    out = []
    lines = src.split("\n")
    i = 0
    depth = 0                      # brace depth, to tell struct members / locals from globals
    in_struct = False
    while i < len(lines):
        raw = lines[i]
        code, comment = strip_comment(raw)
        s = code.strip()
        # R1
        if s.startswith("#version") or re.match(r"layout\s*\(\s*local_size", s):
            out.append("// [R1] " + raw)
            i += 1
            continue
        # R2 / R3 / R4: layout-qualified declarations (may span several lines up to the closing ';')
        if s.startswith("layout"):
            block = [raw]
            while not strip_comment(block[-1])[0].rstrip().endswith(";") or \
                    ("{" in "".join(block) and "}" not in "".join(block)):
                i += 1
                block.append(lines[i])
            text = " ".join(strip_comment(b)[0].strip() for b in block)
            m = re.match(r"layout\s*\([^)]*\)\s*readonly buffer \w+\s*\{\s*(\w+)\s+(\w+)\[\];\s*\};", text)
            if m:
                out.append("static const %s* %s = nullptr;  // [R2]" % (m.group(1), m.group(2)))
            elif "push_constant" in text:
                m = re.match(r"layout\s*\(push_constant\)\s*uniform\s+(\w+)\s*\{(.*)\}\s*(\w+);", text)
                body = m.group(2).strip()
                out.append("static struct %s { %s } %s;  // [R3]" % (m.group(1), body, m.group(3)))
            elif "image2D" in text:
                m = re.search(r"uniform image2D (\w+);", text)
                out.append("static image2D %s;  // [R4]" % m.group(1))
            else:
                out.append("// [R4] dropped (never referenced): " + text)
            i += 1
            continue

        line = code
        # R9
        line = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void glsl_main()", line)
        # R10
        if depth == 0 and re.match(r"\s*const float ", line):
            line = "static " + line.lstrip()
        # R6
        line = re.sub(r"\binout\s+(\w+)\s+(\w+)\s*\[(\d+)\]", r"\1 \2[\3]", line)
        line = re.sub(r"\binout\s+(\w+)\s+(\w+)", r"\1& \2", line)
        # R5
        line = re.sub(r"\b(\w+)\[(\d+)\]\s+(\w+)\s*(;|=)", r"\1 \3[\2]\4", line)
        # R7
        line = FLOAT_LIT.sub(lambda m: m.group(1) + "f", line)
        # struct bookkeeping (for R8)
        m = re.match(r"\s*struct\s+(\w+)\s*\{", line)
        if m:
            STRUCT_TYPES.add(m.group(1))
            in_struct = True
        # R8: plain declarations `T name;` / `T a, b, c;` / `T name[N];` without initialiser
        m = re.match(r"^(\s*)(\w+)\s+([\w\s,\[\]]+);\s*$", line)
        if m and (m.group(2) in BASIC_TYPES or m.group(2) in STRUCT_TYPES) and depth > 0 and m.group(2) != "return":
            names = [n.strip() for n in m.group(3).split(",")]
            line = "%s%s %s;" % (m.group(1), m.group(2), ", ".join(n if ("[" in n and not in_struct) else n + "{}" for n in names))
        depth += line.count("{") - line.count("}")
        if in_struct and depth == 0:
            in_struct = False
        out.append(line + comment)
        i += 1
    return "\n".join(out) + "\n"


def main():
    src_path = sys.argv[1] if len(sys.argv) > 1 else REF_SHADER
    dst_path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                                    "_ref", "raytrace_comp.inc")
    with open(src_path) as f:
        src = f.read()
    os.makedirs(os.path.dirname(dst_path), exist_ok=True)
    with open(dst_path, "w") as f:
        f.write("// GENERATED by oracle/ref_from_glsl.py from %s — the reference's text, do not commit.\n" % src_path)
        f.write(translate(src))
    print(dst_path)


if __name__ == "__main__":
    main()
