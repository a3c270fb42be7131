"""CPU-side checks of bench.py's reference arm and of the oracle/_ref recipe (no GPU)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--steps', '2', '--warmup', '1',
                        '--cpu-size', '96'], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line['impl'] == 'reference' and line['metric'] == 'pds_iterations_per_sec' and line['unit'] == 'iter/s'
    assert line['higher_is_better'] is True and line['value'] > 0 and line['steps'] == 2
    assert line['cpu_baseline']['kind'] in ('reference', 'port') and line['cpu_baseline']['cores'] == 1
    assert line['e2e'] == {'value': line['value'], 'unit': 'iter/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}


@pytest.mark.needs_reference
def test_reference_archive_runs_the_reference_solver(monkeypatch):
    """oracle/make_ref.py packs the reference package; imported from the archive (as on the GPU box) the reference's
    PrimalDualSplitting reproduces the numpy oracle bit for bit."""
    from oracle.make_ref import ARCHIVE, make_ref
    if not make_ref():
        pytest.skip('no reference tree to pack')
    code = (
        "import sys, numpy as np, warnings; warnings.filterwarnings('ignore'); sys.path.insert(0, %r)\n"
        "from oracle import ref_shim, ref_problems as R, np_oracle as O\n"
        "from pycsou_b200.util import synth\n"
        "assert ref_shim.reference_root().endswith('.zip')\n"
        "shape = (40, 56); N = shape[0] * shape[1]\n"
        "_, psf, y = synth.tv_deconv_problem(shape, None, 0.02, 0)\n"
        "_, x, z = R.c3_reference_sample(y, psf, shape, 0.02, 4)\n"
        "import pycsou; assert '.zip' in pycsou.__file__\n"
        "Fo = O.HalfSquaredL2Loss(y, O.Convolve2D(N, psf, shape), G_lipschitz=1.0)\n"
        "xo, zo, _, _ = O.pds_iterate(N, F=Fo, G=O.NonNegativeOrthant(N), H=0.02 * O.TVL21Norm(2 * N, N),\n"
        "    K=O.Gradient(shape, kind='forward'), K_lipschitz=np.sqrt(8.), beta=1.0, max_iter=3, min_iter=3, accuracy_threshold=0)\n"
        "assert np.array_equal(x, xo) and np.array_equal(z, zo)\n" % ROOT)
    env = dict(os.environ, PCS_REFERENCE_ROOT=ARCHIVE)
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
