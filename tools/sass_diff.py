"""Per-function SASS comparison of two `cuobjdump -sass` dumps: proves that a change left the validated kernels untouched.
    cuobjdump -sass lib/libdtc_b200.so > after.sass;  python tools/sass_diff.py before.sass after.sass
Compares instruction text only (addresses, encodings and whitespace are ignored)."""
import re
import sys


def parse(path):
    funcs, cur = {}, None
    for line in open(path, errors="ignore"):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            ins = re.sub(r"/\*.*?\*/", "", line)
            funcs[cur].append(" ".join(ins.split()))
    return funcs


def main() -> int:
    a, b = parse(sys.argv[1]), parse(sys.argv[2])
    changed = [f for f in a if f in b and a[f] != b[f]]
    missing = [f for f in a if f not in b]
    new = [f for f in b if f not in a]
    # a template parameter added to a kernel renames every instantiation: pair renamed functions by identical code
    renamed = []
    for f in list(missing):
        twin = next((g for g in new if b[g] == a[f]), None)
        if twin is not None:
            renamed.append((f, twin))
            missing.remove(f)
            new.remove(twin)
    if renamed:
        print(f"{len(renamed)} functions renamed with identical code")
    print(f"{len(a)} -> {len(b)} functions; changed {len(changed)}, missing {len(missing)}, new {len(new)}")
    for f in changed:
        print(f"  changed  {f[:120]}  ({len(a[f])} -> {len(b[f])} instructions)")
    for f in missing:
        print(f"  missing  {f[:120]}")
    for f in new:
        print(f"  new      {f[:120]}  ({len(b[f])} instructions)")
    return 1 if changed or missing else 0


if __name__ == "__main__":
    sys.exit(main())
