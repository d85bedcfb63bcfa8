"""Developer helper: condensed opcode sequence (loads vs math) of one kernel in the built library."""
import re
import subprocess
import sys

pat = sys.argv[1]
out = subprocess.check_output(["cuobjdump", "-sass", "cassmantle_b200/libcassmantle_b200.so"], text=True)
funcs = re.split(r"\n\s*Function : ", out)
for f in funcs[1:]:
    name = f.split("\n", 1)[0]
    if pat not in name:
        continue
    ops = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_.]+)", f)
    seq, last, cnt = [], None, 0
    for o in ops:
        key = ("LDG128" if o.startswith("LDG") and "128" in o else "LDG" if o.startswith("LDG") else
               "FADD" if o.startswith("FADD") else "SHFL" if o.startswith("SHFL") else
               "LDL" if o.startswith("LDL") else "STL" if o.startswith("STL") else
               "BRA" if o.startswith("BRA") else "CALL" if o.startswith("CALL") else
               "F64" if o[0] == "D" or o.startswith("F2F") or "64" in o and o.startswith("MUFU") else None)
        if key is None:
            continue
        if key == last:
            cnt += 1
        else:
            if last:
                seq.append(f"{last}x{cnt}")
            last, cnt = key, 1
    seq.append(f"{last}x{cnt}")
    print(name[:110])
    print("  total instr", len(ops))
    print("  " + " ".join(seq))
