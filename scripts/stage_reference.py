#!/usr/bin/env python
"""Stage the UNMODIFIED harness modules of an RnB checkout for a run on another machine.

The GPU box has no /root/reference.  This copies the reference's own orchestration files -- the
ones SURVEY.md marks out of scope and reused as they are (benchmark.py, runner.py, control.py,
client.py, rnb_logging.py, selector.py, runner_model.py, video_path_provider.py, arg_utils.py,
batcher.py, utils/class_utils.py) -- byte for byte into baseline/_ref/rnb/, which is git-ignored
(reference sources never enter this repository's history) but travels with a gpurun snapshot.
scripts/run_rnb.py and tests/test_harness.py look there when RNB_REFERENCE is not set.

    python scripts/stage_reference.py [/path/to/rnb]
"""
import hashlib
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = ['benchmark.py', 'runner.py', 'control.py', 'client.py', 'rnb_logging.py', 'selector.py',
         'runner_model.py', 'video_path_provider.py', 'arg_utils.py', 'batcher.py',
         'utils/class_utils.py', 'LICENSE']


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else os.environ.get('RNB_REFERENCE', '/root/reference')
    dst = os.path.join(ROOT, 'baseline', '_ref', 'rnb')
    if not os.path.isdir(src):
        print('no RnB checkout at %s; nothing staged' % src)
        return 1
    for rel in FILES:
        s, d = os.path.join(src, rel), os.path.join(dst, rel)
        os.makedirs(os.path.dirname(d), exist_ok=True)
        shutil.copyfile(s, d)
        print('%s  %s' % (hashlib.sha256(open(d, 'rb').read()).hexdigest()[:16], rel))
    print('staged %d files from %s into %s' % (len(FILES), src, dst))
    return 0


if __name__ == '__main__':
    sys.exit(main())
