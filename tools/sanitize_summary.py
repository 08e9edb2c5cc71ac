"""Condense compute-sanitizer logs into one table (profiles/r2_sanitize_summary.txt)."""
import re
import sys

rows = []
for path in sys.argv[1:]:
    txt = open(path, errors="replace").read()
    m = re.findall(r"(ERROR SUMMARY: \d+ errors?|RACECHECK SUMMARY: [^\n]*)", txt)
    ok = "case ok" in txt
    rows.append((path.split("/")[-1], "ran to completion" if ok else "DID NOT FINISH", "; ".join(m) or "no summary line"))
w = max(len(r[0]) for r in rows) if rows else 10
for r in rows:
    print(f"{r[0]:<{w}}  {r[1]:<18}  {r[2]}")
