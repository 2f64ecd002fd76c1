#!/usr/bin/env python
"""Static SASS opcode table of one kernel of a built object / library (no GPU needed):

  python tools/sass_opcodes.py cudapathtracer_b200/libcupt.so 'k_extendILi0ELb0ELb1E' [out.md]

Prints opcode counts of the first function whose mangled name contains the pattern, and the mnemonics that matter for
the design statements (FFMA2 / FADD2 = packed f32x2, CALL / STL = division slow path and spills, LDG.E.128 = node fetches)."""
import collections
import re
import subprocess
import sys


def functions(path):
    out = subprocess.run(["cuobjdump", "-sass", path], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    name, body = None, []
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            if name:
                yield name, body
            name, body = m.group(1), []
        elif name:
            body.append(line)
    if name:
        yield name, body


def main():
    path, pat = sys.argv[1], sys.argv[2]
    for name, body in functions(path):
        if pat in name:
            break
    else:
        raise SystemExit(f"no function matching {pat}")
    ops, full = collections.Counter(), collections.Counter()
    for line in body:
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_x]+)*)", line)
        if m:
            ops[m.group(1)] += 1
            full[m.group(1) + m.group(2)] += 1
    total = sum(ops.values())
    lines = [f"static SASS of `{name}` in `{path}`: {total} instructions", "", "| opcode | count |", "|---|---:|"]
    lines += [f"| {k} | {v} |" for k, v in ops.most_common(30)]
    keys = ("FFMA2", "FADD2", "FMUL2", "CALL", "STL", "LDL", "FCHK")
    lines += ["", "design markers: " + ", ".join(f"{k} {ops.get(k, 0)}" for k in keys)
              + f", LDG.E.128* {sum(v for k, v in full.items() if k.startswith('LDG.E.128'))}"]
    text = "\n".join(lines) + "\n"
    print(text)
    if len(sys.argv) > 3:
        open(sys.argv[3], "w").write(text)


if __name__ == "__main__":
    main()
