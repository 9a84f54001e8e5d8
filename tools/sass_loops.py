"""Lists the backward-branch loops of a kernel's SASS with their memory instructions (spills / generic loads in
hot loops).  usage: python tools/sass_loops.py <lib.so> <function-substring> [max_len]"""
import re
import subprocess
import sys


def main(lib, func, max_len=400):
    txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    cur, ins = None, {}
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1); ins[cur] = []; continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and cur:
            ins[cur].append((int(m.group(1), 16), m.group(2).strip()))
    for name, code in ins.items():
        if func not in name:
            continue
        addr = {a: i for i, (a, _) in enumerate(code)}
        print(f"== {name}: {len(code)} instructions")
        for i, (a, t) in enumerate(code):
            m = re.search(r"BRA\S*\s+(?:\S+,\s*)*`?\(?0x([0-9a-f]+)", t)
            if not m:
                continue
            tgt = int(m.group(1), 16)
            if tgt in addr and addr[tgt] < i and i - addr[tgt] <= max_len:
                body = code[addr[tgt]:i + 1]
                ops = [x.split()[0] if not x.startswith("@") else x.split()[1] for _, x in body]
                mem = [x for _, x in body if re.search(r"\b(LDL|STL|LD\.E|ST\.E|LDG|STG|LDC)\b", x.replace("LD.E", "LD.E"))]
                nf = sum(o.startswith(("DFMA", "DADD", "DMUL", "DSETP")) for o in ops)
                print(f"  loop [{addr[tgt]}-{i}] len {len(body)}  fp64 {nf}  FSEL {sum(o.startswith('FSEL') for o in ops)} "
                      f"SHFL {sum(o.startswith('SHFL') for o in ops)} LDS {sum(o.startswith('LDS') for o in ops)}")
                for x in mem:
                    print("      mem:", x[:90])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 400)
