"""Per-kernel counts of the SASS mnemonics that prove the hardware path (tcgen05 / TMA / mbarrier / cluster), from
`cuobjdump -sass de_i2i_gan_b200/libdei2i.so`.  Runs in the build container (no GPU):  python tools/sass_summary.py > profiles/<round>_sass_mnemonics.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ['UTCHMMA', 'UTMALDG', 'UTMASTG', 'UTCBAR', 'UTCATOMSWS', 'LDTM', 'SYNCS', 'UCGABAR', 'ELECT', 'LDGSTS', 'HMMA', 'R2UR']


def main():
    sass = subprocess.run(['cuobjdump', '-sass', os.path.join(ROOT, 'de_i2i_gan_b200', 'libdei2i.so')], capture_output=True, text=True, check=True).stdout
    cur, cnt = None, collections.OrderedDict()
    for line in sass.splitlines():
        m = re.match(r'\s*Function : (\S+)', line)
        if m:
            cur = m.group(1)
            cnt[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.search(r'/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*)', line)
        if m:
            op = m.group(1)
            cnt[cur]['_total'] += 1
            for k in KEYS:
                if op.startswith(k):
                    cnt[cur][k] += 1
    names = subprocess.run(['c++filt'], input='\n'.join(cnt), capture_output=True, text=True).stdout.split('\n')
    print("cuobjdump -sass de_i2i_gan_b200/libdei2i.so (sm_100a): instruction counts per kernel")
    print("UTCHMMA = tcgen05.mma (kind::f16 / tf32), UTMALDG = TMA load (tiled / im2col), UTMASTG = TMA store, UTCBAR = tcgen05.commit,")
    print("LDTM = tcgen05.ld, UTCATOMSWS = tcgen05.alloc / dealloc, SYNCS = mbarrier, UCGABAR = barrier.cluster, ELECT = elect.sync,")
    print("LDGSTS = cp.async (gather kernels), HMMA = mma.sync (only the 8 / 16-channel skinny kernels: < 1 % of the FLOPs, HBM- or latency-bound, DESIGN.md section 4), R2UR = register -> uniform register moves\n")
    for (k, c), n in zip(cnt.items(), names):
        n = re.sub(r'\((?:CUtensorMap|dei2i|float|__nv|void|int|unsigned|long|char|bool).*', '', n).replace('void ', '').replace('dei2i::', '')
        tags = ' '.join(f'{q}={c[q]}' for q in KEYS if c[q])
        print(f'{n[:72]:72s} instrs={c["_total"]:6d}  {tags}')


if __name__ == '__main__':
    main()
