"""Markdown rows of BASELINE.md section 4 from a bench.py JSON line (and optionally an N > 1 line for the
train step).  Usage: bench_table.py profiles/<bench>.json [profiles/<bench_n2>.json]"""
import json
import sys


def load(p):
    return json.loads([x for x in open(p).read().splitlines() if x.startswith('{')][-1])


def main():
    d = load(sys.argv[1])
    n2 = load(sys.argv[2]) if len(sys.argv) > 2 else None
    cb = d.get('cpu_baseline') or {}
    rows = []
    r, r2 = d['roofline'], d.get('roofline_pass2', {})
    rows.append(f"| encoder cfg3 (headline, 160 M events, binned path) | {d['value'] / 1e3:.1f} Gev/s, {d['ms_per_step']:.3f} ms/step | "
                f"step {r['step_frac']:.3f}; {r['kernel']} {r['frac']:.3f} ({r['kernel_ms'] * 1e3:.0f} us); "
                f"{r2.get('kernel', '-')} {r2.get('frac', 0):.3f} ({r2.get('kernel_ms', 0) * 1e3:.0f} us) | "
                f"{cb.get('value', 0):.1f} Mev/s ({cb.get('cores')} threads, torch port, 8 of 32 samples) |")
    w = d['workloads']
    for name in ('cfg1', 'cfg2', 'cfg2_contended', 'cfg4'):
        v = w[name]
        ro = v['roofline']
        extra = f"; RED rate {v['red_rate_frac_of_measured_ceiling']:.2f} of the measured 190 G/s ceiling" if 'red_rate_frac_of_measured_ceiling' in v else ''
        p2 = v.get('roofline_pass2')
        extra += f"; {p2['kernel']} {p2['frac']:.3f} ({p2['kernel_ms'] * 1e3:.0f} us)" if p2 else ''
        rows.append(f"| encoder {name} ({v['path'].split(' ')[0]}) | {v['mevents_per_s'] / 1e3:.1f} Gev/s, {v['ms_per_step'] * 1e3:.1f} us/step "
                    f"(warm graph {v['warm_graph_ms_per_step'] * 1e3:.1f} us) | step {v['hbm_frac_step']:.3f}; {ro['kernel']} {ro['frac']:.3f}"
                    f" ({ro['kernel_ms'] * 1e3:.0f} us){extra} | - |")
    for name in ('cfg2_compact', 'cfg3_compact'):
        v = w[name]
        rows.append(f"| encoder {name[:4]}, storage form (9 B/event, {v['path']}) | {v['mevents_per_s'] / 1e3:.1f} Gev/s, "
                    f"{v['ms_per_step'] * 1e3:.1f} us/step | step {v['hbm_frac_step']:.3f} of its own 9 B/event roofline | - |")
    e = d['e2e']
    rows.append(f"| e2e cfg3 through dv.quantize (pinned int64 host tensors, H2D + D2H inside) | {e['value']:.0f} Mev/s; storage form "
                f"{e['compact_storage_form']['value']:.0f} Mev/s | PCIe bound ({e['h2d_bytes_per_step'] / 1e9:.2f} GB / "
                f"{e['compact_storage_form']['h2d_bytes_per_step'] / 1e9:.2f} GB per step) | reference arm: see `--impl reference` line |")
    wl = d['warp_loss']
    c = wl['c_abi']
    cpu = wl.get('cpu_baseline', {})
    rows.append(f"| warp loss cfg3 (B=32, 256x256), C ABI | {c['samples_per_s'] / 1e3:.0f} k samples/s, {c['ms_per_step'] * 1e3:.1f} us/step "
                f"(warm graph {c['warm_graph_ms_per_step'] * 1e3:.1f}) | step {c['hbm_frac_step']:.3f} (SURVEY bytes); fused sweep "
                f"{wl['roofline']['frac']:.3f} of HBM at 24 B/px ({wl['roofline']['kernel_ms'] * 1e3:.1f} us) | "
                f"{cpu.get('all_threads', {}).get('value', 0):.0f} samples/s ({cpu.get('all_threads', {}).get('cores')} threads), "
                f"{cpu.get('one_thread', {}).get('value', 0):.0f} (1 thread), oracle port |")
    rows.append(f"| warp loss cfg3 through training.combined_loss + backward (Python API) | {wl['samples_per_s'] / 1e3:.0f} k samples/s, "
                f"{wl['ms_per_step'] * 1e3:.0f} us/step | {wl['python_over_c_abi']:.1f}x the C-ABI step (host bound: autograd engine + Function) | - |")
    for name in ('cfg1', 'cfg2', 'cfg4'):
        v = w[name + '_loss']
        c = v['c_abi']
        rows.append(f"| warp loss {name}, C ABI | {c['samples_per_s'] / 1e3:.0f} k samples/s, {c['ms_per_step'] * 1e3:.1f} us/step | "
                    f"step {c['hbm_frac_step']:.3f}; fused sweep {v['roofline']['frac']:.3f} ({v['roofline']['kernel_ms'] * 1e3:.1f} us) | - |")
    m = w.get('map_events')
    if m:
        rows.append(f"| map_events (1 M DAVIS events, 17.5 deg LUT) | {m['gpu_mevents_per_s'] / 1e3:.1f} Gev/s ({m['gpu_ms'] * 1e3:.0f} us) | - | "
                    f"{m['cpu_baseline']['value']:.1f} Mev/s, C++ restatement, 1 thread |")
    cd = w.get('codec_cfg2')
    if cd:
        rows.append(f"| batch codec cfg2 (decode / encode, 45 B/event) | {cd['decode_mevents_per_s'] / 1e3:.1f} / {cd['encode_mevents_per_s'] / 1e3:.1f} Gev/s | "
                    f"{cd['decode_hbm_frac']:.2f} / {cd['encode_hbm_frac']:.2f} | - |")
    for k, v in w.items():
        if 'voxel' in k:
            rows.append(f"| EXTENSION (not in the reference, parity unpinned): voxel grid {k} ({v['path']}) | {v['mevents_per_s'] / 1e3:.1f} Gev/s, "
                        f"{v['ms_per_step'] * 1e3:.1f} us/step | step {v['hbm_frac_step']:.3f} (36 N + 4 C B H W bytes) | "
                        f"{v['cpu_baseline']['value']:.1f} Mev/s, numpy oracle, 1 thread |")
    t = d['train_step']
    line = f"| EV-FlowNet train step, 32 samples/GPU | 1 GPU: {t['value']:.0f} samples/s ({t['ms_per_step']:.1f} ms)"
    chk = '-'
    if n2:
        t2 = n2['train_step']
        line += f"; {n2['n_gpus']} GPUs: {t2['value']:.0f} (collate_wrapper(encoded=True) feed), {t2['int64_feed']['value']:.0f} (int64 feed)"
        dc = t2.get('dp_check') or {}
        chk = f"dp_check: allreduced gradient vs mean of rank gradients, max abs err {dc.get('max_abs_err')}"
    rows.append(line + f" | not roofline graded (cuDNN bound) | {chk} |")
    print('\n'.join(rows))


if __name__ == '__main__':
    main()
