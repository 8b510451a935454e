"""Reference checker recipe — TEST / BASELINE INFRASTRUCTURE, run in the BUILD CONTAINER.

    python oracle/make_ref.py        # /root/reference/dgmc  ->  oracle/_ref/dgmc  (git-ignored)

The reference is pure Python, so there is nothing to compile: its package directory is copied
UNMODIFIED into `oracle/_ref/` (listed in .gitignore, not in .gpurunignore: it travels to the GPU box
with the snapshot like a built .so, and never enters the repository's history).  `bench.py --impl
reference` and the `cpu_baseline` leg import it from there, through the third-party stand-ins in
`oracle/shims/` (torch_geometric / torch_scatter are not installable offline), and time the
reference's own `DGMC.forward / loss / backward` on the host cores.  `__graft_entry__.build()` calls
this when /root/reference is present.  A digest of the copied files is written beside them so that
the bench line can say what it ran."""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get('DGMC_REFERENCE', '/root/reference')
DST = os.path.join(HERE, '_ref')


def make():
    src = os.path.join(SRC, 'dgmc')
    if not os.path.isdir(src):
        return None
    dst = os.path.join(DST, 'dgmc')
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    os.makedirs(DST, exist_ok=True)
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns('__pycache__', '*.pyc'))
    h = hashlib.sha256()
    for root, _, files in sorted(os.walk(dst)):
        for f in sorted(files):
            h.update(open(os.path.join(root, f), 'rb').read())
    with open(os.path.join(DST, 'DIGEST'), 'w') as fh:
        fh.write(h.hexdigest() + '\n')
    return dst


if __name__ == '__main__':
    out = make()
    print(out or ('no reference at %s' % SRC))
    sys.exit(0 if out else 1)
