"""Print per-kernel durations (us) from an ncu --csv --metrics gpu__time_duration.sum log (development helper)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
for r in rows:
    if len(r) > 10 and r[0].isdigit() and "gpu__time_duration" in r[-3]:
        print(f"{float(r[-1].replace(',', '')) / 1e3:9.1f} us  {r[4][:60]}")
