"""Per-SASS-instruction executed counts and stall samples from a .ncu-rep (page source), grouped into
regions by executed-count changes. Usage: ncu_sass.py rep [min_count]"""
import csv, subprocess, sys, io
def main(path, full=False):
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ia, isrc, isamp, iex = hdr.index('Address'), hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
    tot = sum(int(r[iex]) for r in rows[2:] if len(r) > iex)
    tots = sum(int(r[isamp]) for r in rows[2:] if len(r) > iex)
    print('total inst', tot, 'samples', tots)
    for n, r in enumerate(rows[2:]):
        if len(r) <= iex: continue
        print('%4d %10d %6d  %s' % (n, int(r[iex]), int(r[isamp]), r[isrc].strip()[:90]))
if __name__ == '__main__':
    main(sys.argv[1])
