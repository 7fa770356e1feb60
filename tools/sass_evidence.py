"""Regenerate profiles/sass/SASS_EVIDENCE.md from the built library (cuobjdump -sass)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "mpi_jordan_crazy_acceleration_b200", "lib", "libjordan_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
keys = ['DMMA', 'UTCIMMA', 'UTCHMMA', 'LDTM', 'STTM', 'UTMALDG', 'UTMAREDG', 'UTMASTG', 'UBLKCP', 'UBLKRED', 'SYNCS',
        'UCGABAR', 'REDUX', 'STG.E.128.STRONG.SYS', 'STG.E.64.STRONG.SYS', 'REDG.E.ADD.STRONG.SYS', 'LDG.E.STRONG.SYS',
        'UTCBAR', 'UTCATOMSWS', 'LDGSTS', 'DFMA', 'MUFU.RCP64H', 'SHFL', 'BAR.SYNC']
out = ['# SASS evidence per kernel (`cuobjdump -sass libjordan_b200.so`, sm_100a)\n', '| kernel | SASS instructions | evidence |', '|---|---|---|']
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n', 1)[0].strip()
    dem = subprocess.run(['c++filt', name], capture_output=True, text=True).stdout.strip()
    short = dem.replace('(anonymous namespace)::', '').replace('void ', '').split('(')[0]
    ops = re.findall(r'/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)', f)
    cnt = collections.Counter(k for o in ops for k in keys if o.startswith(k))
    out.append(f"| `{short}` | {len(ops)} | " + ', '.join(f'{k} ×{v}' for k, v in cnt.most_common()) + " |")
print("\n".join(out))
