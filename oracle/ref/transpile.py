#!/usr/bin/env python
"""ORACLE - TEST INFRASTRUCTURE ONLY.

Rewrites the reference's GLSL shader sources, read where they lie under /root/reference/src, into files that
g++ accepts, under oracle/_ref/gen/ (git-ignored: nothing of the reference enters the repository).  The shaders
are nearly C++ already; what is rewritten, mechanically and line by line:

  * `#version` / `#extension` lines and `layout(...)` qualifiers are dropped;
  * interface blocks become plain globals: `uniform X { T a; };` -> `T a;`,  `readonly buffer X { T a[]; };` ->
    `T* a;`,  `uniform sampler2DArray S;` -> `sampler2DArray S;`;
  * parameter qualifiers: `out T a` / `inout T a` -> `T& a`,  `in T a` -> `T a`;
  * the swizzles the shaders use (.xyz .xy .yx .yz .xz .rgb) become member calls (.xyz());
  * floating literals get the `f` suffix (GLSL literals are single precision);
  * a vector constructor with more than one Random*() call among its arguments is written with braces: GLSL
    evaluates arguments left to right, C++ only does so inside a braced initialiser list;
  * array constructors `T[n](a, b, ..)` become braced initialisers; stage inputs / outputs (`in vec2 X;`) globals;
  * `void main()` is renamed after the file (basic_trace.glsl -> basic_trace_main).

oracle/ref/ref_glsl.h supplies the GLSL types and built-ins; oracle/ref/ref_api.cpp includes the generated files
inside namespace glsl and drives them.  Usage: python oracle/ref/transpile.py [REFERENCE_SRC] [OUT_DIR]
"""
import os
import re
import sys

FILES = [
    "core/common.glsl.inc", "core/spectrum.glsl.inc", "scene/scene.glsl.inc", "scene/basic_diffuse.glsl.inc",
    "scene/basic_metal.glsl.inc", "scene/basic_translucent.glsl.inc", "integrator/basic.glsl.inc",
    "integrator/basic_trace.glsl", "integrator/basic_scatter.glsl", "integrator/resolve.glsl",
    "application/preview_render.glsl",
]

FLOAT_LITERAL = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][-+]?\d+)?|\d+[eE][-+]?\d+)(?![\w.])")
SWIZZLE = re.compile(r"\.(xyz|xy|yx|yz|xz|rgb)\b(?!\s*\()")
OUT_PARAM = re.compile(r"\b(?:out|inout)\s+(\w+)\s+(\w+)")
IN_PARAM = re.compile(r"\bin\s+(\w+)\s+(\w+)")
CONSTRUCTOR = re.compile(r"\b(?:vec[234])\s*\(")
LAYOUT = re.compile(r"\blayout\s*\([^)]*\)\s*")


def strip_comment(line):
    i = line.find("//")
    return (line, "") if i < 0 else (line[:i], line[i:])


def sequence_constructor_arguments(code):
    """vecN(.. Random*() .., .. Random*() ..) -> vecN{...}: left-to-right evaluation as GLSL specifies."""
    for m in reversed(list(CONSTRUCTOR.finditer(code))):
        depth, j = 0, m.end() - 1
        for j in range(m.end() - 1, len(code)):
            depth += (code[j] == "(") - (code[j] == ")")
            if depth == 0:
                break
        else:
            continue
        if len(re.findall(r"\bRandom\w*\s*\(", code[m.end():j])) > 1:
            code = code[:m.end() - 1] + "{" + code[m.end():j] + "}" + code[j + 1:]
    return code


def rewrite_code(code):
    code = sequence_constructor_arguments(code)
    code = FLOAT_LITERAL.sub(lambda m: m.group(1) + "f", code)
    code = SWIZZLE.sub(lambda m: "." + m.group(1) + "()", code)
    code = OUT_PARAM.sub(r"\1& \2", code)
    code = IN_PARAM.sub(r"\1 \2", code)
    return code


ARRAY_CONSTRUCTOR = re.compile(r"=\s*\w+\s*\[\s*\d*\s*\]\s*\(")


def brace_array_constructors(text):
    """`= vec3[20]( ... )` -> `= { ... }` (may span lines)."""
    while True:
        m = ARRAY_CONSTRUCTOR.search(text)
        if not m:
            return text
        depth, j = 0, m.end() - 1
        for j in range(m.end() - 1, len(text)):
            depth += (text[j] == "(") - (text[j] == ")")
            if depth == 0:
                break
        text = text[:m.start()] + "= {" + text[m.end():j] + "}" + text[j + 1:]


def transpile(text, stem):
    text = brace_array_constructors(text)
    out = []
    lines = text.split("\n")
    i = 0
    in_block_comment = False
    while i < len(lines):
        line = lines[i]
        i += 1
        if in_block_comment:
            out.append(line)
            if "*/" in line:
                in_block_comment = False
            continue
        if line.lstrip().startswith("/*") and "*/" not in line:
            in_block_comment = True
            out.append(line)
            continue
        s = line.strip()
        if s.startswith("#version") or s.startswith("#extension"):
            out.append("// " + s)
            continue
        if s.startswith("#"):
            out.append(line)                    # #include / #define / #if: kept (includes resolve inside gen/)
            continue
        code, comment = strip_comment(line)
        if LAYOUT.search(code):
            code = LAYOUT.sub("", code)
            # the qualifier may be followed on this or the next lines by:  in;  |  uniform T name;  |  [readonly] [uniform|buffer] Name { members };
            rest = code.strip()
            while rest == "" and i < len(lines):
                nxt, _ = strip_comment(lines[i])
                i += 1
                rest = nxt.strip()
            if rest == "in;":
                out.append("// (work-group size)")
                continue
            m = re.match(r"^(?:readonly\s+|writeonly\s+)?(?:uniform|buffer)\s+(\w+)\s*$", rest)
            if m:                                # interface block: members become globals
                body = []
                while i < len(lines):
                    b, _ = strip_comment(lines[i])
                    i += 1
                    if b.strip().startswith("}"):
                        break
                    if b.strip() not in ("{", ""):
                        body.append(b.strip())
                out.append("// interface block " + m.group(1))
                for member in body:
                    mm = re.match(r"^(\w+)\s+(\w+)\s*\[\s*\]\s*;$", member)
                    out.append("%s* %s;" % (mm.group(1), mm.group(2)) if mm else rewrite_code(member))
                continue
            m = re.match(r"^(?:uniform|in|out)\s+(\w+)\s+(\w+)\s*;$", rest)
            if m:
                out.append("%s %s;" % (m.group(1), m.group(2)))
                continue
            raise SystemExit("transpile: unhandled layout declaration in %s: %r" % (stem, rest))
        code = rewrite_code(code)
        code = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void %s_main()" % stem, code)
        out.append(code + comment)
    return "\n".join(out)


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/src"
    dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "_ref", "gen")
    for rel in FILES:
        with open(os.path.join(src, rel)) as f:
            text = f.read()
        stem = os.path.basename(rel).split(".")[0]
        path = os.path.join(dst, rel)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w") as f:
            f.write("// GENERATED by oracle/ref/transpile.py from %s - do not commit\n" % rel)
            f.write(transpile(text, stem))
    print("transpiled %d files into %s" % (len(FILES), os.path.normpath(dst)))


if __name__ == "__main__":
    main()
