#!/usr/bin/env python
"""Bucket the per-line output of ncu_src_top.py by phase of export_world (line ranges found by marker text)."""
import re, sys
src = open('/root/repo/madrona-bots_b200/csrc/sim_kernels.cu').read().split('\n')
def find(txt): return next(i + 1 for i, l in enumerate(src) if txt in l)
marks = [(find('template <typename V, int PARTS'), 'carry'), (find('One world, one warp.'), 'per-agent export'),
         (find('the sort: the wide columns'), 'carry'), (find('if (!with_sensors) return;'), 'best init'),
         (find('auto test = '), 'exact test'), (find('auto run_tests'), 'exact test'), (find('auto expand = '), 'expand'),
         (find('auto spans = '), 'cull stage 2'), (find('int cqn = 0;'), 'cull stage 1'),
         (find('for (int idx = lane; idx < m * 8'), 'sensor output'), (find('skip: FAST -> worlds'), 'other')]
psrc = open('/root/repo/madrona-bots_b200/csrc/sim_prims.cuh').read().split('\n')
def pfind(txt): return next(i + 1 for i, l in enumerate(psrc) if txt in l)
pm = [(0, 'per-agent export'), (pfind('MBP_HD Ray make_ray'), 'exact test'), (pfind('conservative candidate rays'), 'cull stage 2'),
      (pfind('MBP_HD Dir cull_heading'), 'per-agent export'), (pfind('MBP_HD float cull_atan2'), 'cull stage 2'),
      (pfind('MBP_HD bool pair_maybe'), 'cull stage 1'), (pfind('MBP_HD bool pair_candidates'), 'other')]
def bucket(f, l):
    if f == 'sim_kernels.cu':
        b = 'other'
        for s, n in marks:
            if l >= s: b = n
        return b
    if f == 'sim_prims.cuh':
        b = 'other'
        for s, n in pm:
            if l >= s: b = n
        return b
    if f == 'mbots_detmath.h': return 'detmath fma (cull+test)' if l < 40 else 'detmath sincos/atan2'
    return f
b = {}
for ln in open(sys.argv[1]):
    m = re.match(r'\s*([\d.]+)% smp\s+([\d.]+)% ins\s+(\S+):(\d+)', ln)
    if not m: continue
    k = bucket(m[3], int(m[4]))
    a = b.setdefault(k, [0, 0]); a[0] += float(m[1]); a[1] += float(m[2])
for k, v in sorted(b.items(), key=lambda x: -x[1][1]): print(f"{k:30s} smp {v[0]:5.1f}%  ins {v[1]:5.1f}%")
