"""Which kernels changed since a given commit?  Compiles csrc/*.cu of <rev> and of the working tree for sm_100a and
compares the SASS of every kernel (addresses masked).  Used to prove that adding an opt-in kernel left the binaries of
the hardware-validated kernels untouched:   python tools/sass_diff.py <rev>      (no GPU needed)"""
import os
import re
import subprocess
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-I" + os.path.join(REPO, "include")]


def kernels(obj):
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout
    d, cur = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            d[cur] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if m and cur:
            d[cur].append(re.sub(r"0x[0-9a-f]+", "0x", m.group(1)).strip())
    return d


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return {n: re.sub(r"\(.*", "", o) for n, o in zip(names, out)}


def main():
    rev = sys.argv[1]
    with tempfile.TemporaryDirectory() as tmp:
        old_src = os.path.join(tmp, "old")
        os.makedirs(old_src)
        files = subprocess.run(["git", "ls-tree", "--name-only", rev, "csrc/"], cwd=REPO, capture_output=True,
                               text=True, check=True).stdout.split()
        for f in files:
            with open(os.path.join(old_src, os.path.basename(f)), "w") as fh:
                fh.write(subprocess.run(["git", "show", f"{rev}:{f}"], cwd=REPO, capture_output=True, text=True,
                                        check=True).stdout)
        changed = 0
        for f in sorted(x for x in os.listdir(os.path.join(REPO, "csrc")) if x.endswith(".cu")):
            new_o, old_o = os.path.join(tmp, f + ".new.o"), os.path.join(tmp, f + ".old.o")
            subprocess.run(["nvcc"] + FLAGS + ["-I" + os.path.join(REPO, "csrc"), "-c",
                                               os.path.join(REPO, "csrc", f), "-o", new_o], check=True)
            new = kernels(new_o)
            if not os.path.exists(os.path.join(old_src, f)):
                print(f"{f}: new file, {len(new)} kernels")
                continue
            subprocess.run(["nvcc"] + FLAGS + ["-I" + old_src, "-c", os.path.join(old_src, f), "-o", old_o],
                           check=True)
            old = kernels(old_o)
            dn, do = demangle(list(new)), demangle(list(old))
            by_name = {}
            for k, v in new.items():
                by_name.setdefault(dn[k], []).append(v)
            same = diff = 0
            for k, v in old.items():
                # a kernel that gained a defaulted template parameter keeps its code but not its name: accept any new
                # kernel whose demangled name starts with the old one's (minus the closing '>') and has identical SASS
                stem = do[k].rstrip(">")
                cands = [vv for name, vs in by_name.items() if name.startswith(stem) for vv in vs]
                if v in cands:
                    same += 1
                else:
                    diff += 1
                    print(f"  CHANGED: {do[k]}")
            changed += diff
            print(f"{f}: {same} kernels identical, {diff} changed, {len(new) - same - diff} new")
        return 1 if changed else 0


if __name__ == "__main__":
    sys.exit(main())
