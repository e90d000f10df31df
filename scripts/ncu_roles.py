"""Per-role stall summary of a warp-specialised kernel from an ncu report (development aid).
usage: python scripts/ncu_roles.py report.ncu-rep  [min_samples]
Roles are recognised by marker opcodes in address order (HMMA = x producer, UTCHMMA = MMA issuer,
STTM = convert, LDTM = drain); every instruction with >= min_samples samples is listed."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
mins = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
data = [r for r in rows[2:] if len(r) == len(h) and r != h]
iS, iSrc, iE = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
st = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
MARK = [("HMMA.16816", "producer"), ("UTCHMMA", "issuer"), ("STTM", "convert"), ("LDTM", "drain")]
role, roles = "prologue", []
first_seen = {}
for i, r in enumerate(data):
    for m, name in MARK:
        if m in r[iSrc] and name not in first_seen:
            first_seen[name] = i
# a role's code starts a little before its first marker; use the preceding USETMAXREG / BSSY as hint: simply cut at markers
bounds = sorted((i, n) for n, i in first_seen.items())
def role_of(i):
    cur = "prologue"
    for b, n in bounds:
        if i >= b - 400 and n != cur:
            pass
    # nearest marker at or after a lookback window
    cur = "prologue"
    for b, n in bounds:
        if i >= b - 150:
            cur = n
    return cur
tot = {}
for i, r in enumerate(data):
    tot[role_of(i)] = tot.get(role_of(i), 0) + int(r[iS])
print("samples per role:", tot, "total", sum(tot.values()))
for i, r in enumerate(data):
    if int(r[iS]) >= mins:
        s = sorted(((h[j][6:], int(r[j])) for j in st if int(r[j]) > 0), key=lambda kv: -kv[1])[:2]
        print(f"{i:6d} {role_of(i):9s} {r[iS]:>5} {r[iE]:>8}  {r[iSrc].strip()[:64]:<64} {s}")
