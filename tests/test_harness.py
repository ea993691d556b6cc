"""The step classes inside RnB itself: the reference's unmodified benchmark.py / client.py /
runner.py / control.py drive `rnb_b200.steps.*` from the pipeline files in config/, exactly as a user
who switches the `model` paths of a pipeline file would run them (north star: "drops into the existing
client/runner/queue benchmark").

The reference tree is not part of this repository.  The tests look for it in $RNB_REFERENCE, then in
baseline/_ref/rnb (filled by scripts/stage_reference.py, git-ignored), and skip when neither exists.
"""
import glob
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'scripts'))


def _reference():
    import run_rnb
    return run_rnb.find_reference()


def _run(tmp_path, config, videos, extra=(), timeout=900):
    logdir = str(tmp_path / 'rnb')
    cmd = [sys.executable, os.path.join(ROOT, 'scripts', 'run_rnb.py'), '--logdir', logdir,
           *extra, '-c', os.path.join(ROOT, 'config', config), '-v', str(videos), '-mi', '0']
    env = dict(os.environ)
    env.setdefault('CUDA_VISIBLE_DEVICES', '0')
    env['RNB_B200_ALLOW_BUSY_GPUS'] = '1'   # this pytest process holds a context on the GPU
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                          timeout=timeout, env=env)
    assert proc.returncode == 0, proc.stdout[-3000:]
    jobs = glob.glob(os.path.join(logdir, 'logs', '*'))
    assert len(jobs) == 1, proc.stdout[-2000:]
    import latency_percentiles
    return latency_percentiles.parse_job(jobs[0], skip=0), proc.stdout


@pytest.mark.parametrize('wire', ['fp32', 'uint8'])
def test_whole_pipeline_under_the_reference_harness(lib, tmp_path, wire):
    """config/r2p1d-whole.json: bulk client -> filename queue -> loader step -> shared tensors ->
    runner step (layers 1-5), every process and queue the reference's own."""
    if _reference() is None:
        pytest.skip('no RnB checkout (RNB_REFERENCE or baseline/_ref/rnb)')
    out, log = _run(tmp_path, 'r2p1d-whole.json', 60, extra=('--wire', wire))
    assert out['videos_logged'] >= 60
    assert out['termination_flag'] == 0           # TARGET_NUM_VIDEOS_REACHED, control.py:13
    assert out['keys'][0] == 'enqueue_filename' and out['keys'][-1] == 'inference1_finish'
    assert out['latency_ms']['p50'] > 0 and out['videos_per_s'] > 0
    assert 'Finished processing 60 videos' in log


def test_whole_pipeline_with_the_event_fenced_worker_loop(lib, tmp_path):
    """Same run with rnb_b200.runtime.runner in place of the reference's runner.py: uint8 wire,
    ready / free CUDA events instead of copy_ + stream.synchronize(); client, queues, control plane
    and logs still the reference's."""
    if _reference() is None:
        pytest.skip('no RnB checkout (RNB_REFERENCE or baseline/_ref/rnb)')
    out, log = _run(tmp_path, 'r2p1d-whole.json', 60, extra=('--wire', 'uint8', '--runtime', 'b200'))
    assert out['videos_logged'] >= 60 and out['termination_flag'] == 0
    assert 'Finished processing 60 videos' in log


def test_single_step_pipeline_under_the_reference_harness(lib, tmp_path):
    """config/r2p1d-nopipeline.json on one GPU: R2P1DSingleStep as the only (and final) step."""
    if _reference() is None:
        pytest.skip('no RnB checkout (RNB_REFERENCE or baseline/_ref/rnb)')
    cfg = json.load(open(os.path.join(ROOT, 'config', 'r2p1d-nopipeline.json')))
    cfg['pipeline'][0]['queue_groups'][0]['gpus'] = [0]
    path = tmp_path / 'nopipeline-1gpu.json'
    path.write_text(json.dumps(cfg))
    logdir = str(tmp_path / 'rnb')
    cmd = [sys.executable, os.path.join(ROOT, 'scripts', 'run_rnb.py'), '--logdir', logdir,
           '-c', str(path), '-v', '40', '-mi', '0']
    env = dict(os.environ)
    env.setdefault('CUDA_VISIBLE_DEVICES', '0')
    env['RNB_B200_ALLOW_BUSY_GPUS'] = '1'   # this pytest process holds a context on the GPU
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                          timeout=900, env=env)
    assert proc.returncode == 0, proc.stdout[-3000:]
    import latency_percentiles
    out = latency_percentiles.parse_job(glob.glob(os.path.join(logdir, 'logs', '*'))[0], skip=0)
    assert out['videos_logged'] >= 40 and out['termination_flag'] == 0
