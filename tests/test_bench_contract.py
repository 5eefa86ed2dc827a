"""The reference arm of bench.py (`--impl reference`) runs on host cores only, so its JSON contract can be checked
without a GPU: one line, the keys the driver reads, the reference's own algorithm (oracle/dense_port.py) as `kind`."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '1',
                          '--warmup', '3'], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and d['n_gpus'] == 1 and d['steps'] == 1 and d['warmup'] >= 3
    assert d['metric'].startswith('NT-Xent fwd+bwd pairs/s') and d['unit'] == 'pairs/s' and d['higher_is_better'] is True
    assert d['value'] > 0 and abs(d['value'] - 4096 / (d['ms_per_step'] * 1e-3)) <= 1e-6 * d['value']
    assert d['vs_baseline'] is None and d['data'] == 'synthetic' and 'workload' in d['config']
    cb = d['cpu_baseline']
    assert cb['kind'] in ('port', 'reference') and cb['cores'] >= 1 and cb['value'] == d['value'] and cb['sample']
    assert d['e2e'] == {'value': d['value'], 'unit': 'pairs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert d['gpu_launches'] == 0                      # nothing of this library runs on the reference arm
    sb = d['syncbn']
    assert sb['unit'] == 'GB/s' and sb['value'] > 0
