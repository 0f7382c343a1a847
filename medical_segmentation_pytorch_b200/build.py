"""In-tree build of libmsx.so (sm_100a only) with nvcc; no torch headers are involved.

``python medical_segmentation_pytorch_b200/build.py`` or ``__graft_entry__.build()``.
nvcc cross-compiles without a GPU; the .so stays next to this file (git-ignored, but it travels
to the GPU box with the gpurun snapshot).
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, 'libmsx.so')
STAMP = os.path.join(HERE, 'csrc', '.build_stamp')
SOURCES = ['api.cu', 'norm.cu', 'conv_simt.cu', 'conv_tc.cu', 'conv_api.cu', 'loss.cu', 'head.cu', 'optim.cu', 'noalloc.cu']
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17',
              '-Xcompiler', '-fPIC', '--expt-relaxed-constexpr']


def _nvcc():
    for c in (os.environ.get('NVCC'), shutil.which('nvcc'), '/usr/local/cuda/bin/nvcc'):
        if c and os.path.exists(c):
            return c
    raise RuntimeError('nvcc not found')


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _digest():
    h = hashlib.sha256()
    names = sorted(os.listdir(CSRC)) + ['../../include/msx.h']
    for n in names:
        p = os.path.join(CSRC, n)
        if os.path.isfile(p) and (n.endswith(('.cu', '.cuh', '.h'))):
            h.update(n.encode())
            h.update(open(p, 'rb').read())
    h.update(' '.join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_variant(name, defines, verbose=False):
    """Measurement builds: the same sources with extra -D switches into ``lib<name>.so`` next to libmsx.so
    (select it at run time with MSX_LIB=<path>); never the default library."""
    nvcc = _nvcc()
    objdir = os.path.join(HERE, 'build', name)
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for s in _sources():
        obj = os.path.join(objdir, s.replace('.cu', '.o'))
        procs.append((s, obj, subprocess.Popen([nvcc] + NVCC_FLAGS + list(defines) + ['-c', os.path.join(CSRC, s), '-o', obj] + (['-Xptxas', '-v'] if verbose else []),
                                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    for s, obj, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f'nvcc failed on {s}:\n{out}')
        if verbose and out:
            print(out)
        objs.append(obj)
    lib = os.path.join(HERE, f'lib{name}.so')
    r = subprocess.run([nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', lib] + objs, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f'link failed:\n{r.stdout}')
    return lib


def build(force=False, verbose=False):
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP) and open(STAMP).read().strip() == dig:
        return LIB
    nvcc = _nvcc()
    objdir = os.path.join(HERE, 'build')
    os.makedirs(objdir, exist_ok=True)
    procs = []
    for s in _sources():
        obj = os.path.join(objdir, s.replace('.cu', '.o'))
        cmd = [nvcc] + NVCC_FLAGS + ['-c', os.path.join(CSRC, s), '-o', obj]
        if verbose:
            cmd += ['-Xptxas', '-v']
        procs.append((s, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    objs = []
    for s, obj, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f'nvcc failed on {s}:\n{out}')
        if verbose and out:
            print(out)
        objs.append(obj)
    cmd = [nvcc, '-shared', '-gencode', 'arch=compute_100a,code=sm_100a', '-o', LIB] + objs
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f'link failed:\n{r.stdout}')
    with open(STAMP, 'w') as f:
        f.write(dig)
    return LIB


if __name__ == '__main__':
    if '--variant' in sys.argv:          # python build.py --variant msx_alt -DMSX_TC_SMEM_KB=150 ...
        i = sys.argv.index('--variant')
        print(build_variant(sys.argv[i + 1], [a for a in sys.argv[i + 2:] if a.startswith('-D')], verbose='-v' in sys.argv))
    else:
        print(build(force='--force' in sys.argv, verbose='-v' in sys.argv))
