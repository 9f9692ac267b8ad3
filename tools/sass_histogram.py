"""Static SASS opcode counts of the shipped objects: python tools/sass_histogram.py > profiles/r02_sass_histogram.txt
(after python -m drynx_b200.build gpu; cuobjdump -sass on build/*.o, one section per kernel of interest)."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KERNELS = [
    ('kernels_rp_ucode_s4.o', 'k_pair_ucode_s4E', 'k_pair_ucode_s4 (lane-group pairing, static launch; the interpreter loop with every Fp2 operation inlined)'),
    ('kernels_rp_ucode_s8.o', 'k_pair_ucode_s8E', 'k_pair_ucode_s8 (the same with eight program lanes per check)'),
    ('kernels_rp_gtlanes.o', 'k_rp_create_gt_lanes', 'k_rp_create_gt_lanes (GT factors of proof creation on lane groups)'),
    ('kernels_g1.o', 'k_probe_montmul', 'k_probe_montmul (drx_microbench kind 1: the loop body is two calls of the out-of-line fp_mul; its callee is listed inside the kernel)'),
    ('kernels_g1.o', '9k_encryptE', 'k_encrypt'),
    ('kernels_g1.o', 'k_agg_partial', 'k_agg_partial'),
    ('kernels_g1.o', '9k_decryptE', 'k_decrypt (variable-base product by the GLV walk)'),
    ('kernels_rp.o', 'k_rp_verify_hdr', 'k_rp_verify_hdr (two threads per proof)'),
]


def sections(obj):
    txt = subprocess.run(['cuobjdump', '-sass', os.path.join(ROOT, 'build', obj)], capture_output=True, text=True).stdout
    out, name = {}, None
    for line in txt.split('\n'):
        m = re.search(r'Function : (\S+)', line)
        if m:
            name = m.group(1)
            out[name] = []
        elif name and re.search(r'/\*[0-9a-f]{4,6}\*/', line):
            m = re.search(r'/\*[0-9a-f]{4,6}\*/\s+(.*?);', line)
            if m:
                out[name].append(re.sub(r'^@!?U?P\d\s+', '', m.group(1).strip()).split()[0])
    return out


def main():
    print('# cuobjdump -sass opcode counts (static: instructions in the binary, not executed counts) of the shipped objects, end-of-round-2 build.')
    print('# Regenerate: python -m drynx_b200.build gpu && python tools/sass_histogram.py')
    print('# IMAD.WIDE.U32(.X) is the 32x32->64 multiply-accumulate the INT roofline counts (one per MAC32); a Montgomery product is 128 of them.')
    cache = {}
    for obj, key, title in KERNELS:
        if obj not in cache:
            cache[obj] = sections(obj)
        hits = [n for n in cache[obj] if key in n]
        if not hits:
            continue
        ops = cache[obj][sorted(hits, key=len)[0]]
        c = collections.Counter(ops)
        wide = sum(v for k, v in c.items() if k.startswith('IMAD.WIDE.U32'))
        print('\n== %s' % title)
        print('   %d instructions, %d IMAD.WIDE.U32(.X) (%.1f %%)' % (len(ops), wide, 100.0 * wide / max(1, len(ops))))
        for k, v in c.most_common(18):
            print('   %7d %s' % (v, k))


if __name__ == '__main__':
    main()
