"""SASS opcode histogram per kernel of a built library (cuobjdump -sass): python scripts/sass_hist.py lib.so [kernel-substring]"""
import collections
import re
import subprocess
import sys


def hist(lib, filt=None):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    fn = None; res = collections.OrderedDict()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1); res[fn] = collections.Counter(); continue
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", line)
        if m and fn:
            res[fn][m.group(1)] += 1
    return {k: v for k, v in res.items() if not filt or filt in k}


if __name__ == "__main__":
    for k, v in hist(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None).items():
        tot = sum(v.values())
        print("%s  (%d instructions)" % (k, tot))
        print("   " + "  ".join("%s %d" % (o, c) for o, c in v.most_common(28)))
