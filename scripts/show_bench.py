import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, {k: (round(d[k], 3) if isinstance(d[k], float) else d[k]) for k in ['value', 'ms_per_step', 'ms_per_denoise_step', 'tensor_frac_whole_loop', 'dtype'] if k in d})
    if 'e2e' in d: print('  e2e', round(d['e2e']['value'], 1))
    if 'roofline' in d: print('  roofline', round(d['roofline']['achieved'], 1), round(d['roofline']['frac'], 3))
    for k, v in d.get('kernels', {}).items():
        print('    %-38s n=%d  %.4f ms  share %.3f  %s' % (k, v['launches_per_step'], v['ms_per_launch'], v['share'], ('%.0f TF' % v['tflops']) if 'tflops' in v else ''))
