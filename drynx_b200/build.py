"""Builds the native pieces in-tree with nvcc / g++ (no torch extension machinery, no JIT cache):

  drynx_b200/libdrynx_gpu.so      the product: sm_100a kernels + C ABI (include/drynx_gpu.h)
  oracle/_build/liboracle.so      TEST INFRASTRUCTURE: C++ restatement of the reference's CPU algorithms
  tests/hostemu/libhostemu.so     TEST INFRASTRUCTURE: host build of the device headers (carry-flag emulation)

Run:  python -m drynx_b200.build [gpu|oracle|hostemu|all]
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'drynx_b200', 'csrc')
GPU_SO = os.environ.get('DRYNX_LIB', os.path.join(ROOT, 'drynx_b200', 'libdrynx_gpu.so'))
ORACLE_SO = os.path.join(ROOT, 'oracle', '_build', 'liboracle.so')
HOSTEMU_SO = os.path.join(ROOT, 'tests', 'hostemu', 'libhostemu.so')
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-std=c++17', '-Xcompiler', '-fPIC',
              '-Xptxas', '-v'] + (['-DDRX_DEBUG_RP'] if os.environ.get('DRX_DEBUG_RP') else []) + \
    [f for f in os.environ.get('DRX_NVCC_EXTRA', '').split() if f]
GPU_SOURCES = ['kernels_g1.cu', 'kernels_rp.cu', 'kernels_rp_ucode.cu', 'kernels_rp_ucode_s4.cu', 'kernels_rp_gtlanes.cu', 'kernels_rp_ucode_s8.cu', 'kernels_ed.cu', 'capi.cu']


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _headers():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h', '.inc'))] + \
        [os.path.join(ROOT, 'include', 'drynx_gpu.h')]


def _run(cmd, log=None):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log:
        with open(log, 'w') as f:
            f.write(' '.join(cmd) + '\n' + p.stdout)
    if p.returncode != 0:
        raise RuntimeError('build failed: %s\n%s' % (' '.join(cmd), p.stdout[-4000:]))
    return p.stdout


def build_gpu(force=False):
    objdir = os.environ.get('DRX_OBJDIR', os.path.join(ROOT, 'build'))
    os.makedirs(objdir, exist_ok=True)
    hdrs = _headers()
    jobs = []
    for src in GPU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace('.cu', '.o'))
        if force or _newer(o, [s] + hdrs):
            jobs.append((NVCC_FLAGS and [NVCC] + NVCC_FLAGS + ['-I', CSRC, '-c', s, '-o', o], o + '.log'))
    with ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(lambda j: _run(*j), jobs))
    objs = [os.path.join(objdir, s.replace('.cu', '.o')) for s in GPU_SOURCES]
    if force or jobs or _newer(GPU_SO, objs):
        _run([NVCC, '-shared', '-o', GPU_SO] + objs + ['-lcudart'])
    return GPU_SO


def build_oracle(force=False):
    srcs = [os.path.join(ROOT, 'oracle', 'oracle.cpp'), os.path.join(ROOT, 'oracle', 'oracle_ed.cpp')]
    os.makedirs(os.path.dirname(ORACLE_SO), exist_ok=True)
    deps = srcs + [os.path.join(ROOT, 'oracle', f) for f in os.listdir(os.path.join(ROOT, 'oracle'))
                   if f.endswith('.hpp') or f.endswith('.inc')]
    if force or _newer(ORACLE_SO, deps):
        _run(['g++', '-O3', '-march=native', '-std=c++17', '-shared', '-fPIC', '-pthread'] + srcs + ['-o', ORACLE_SO])
    return ORACLE_SO


def build_hostemu(force=False):
    src = os.path.join(ROOT, 'tests', 'hostemu', 'hostemu.cpp')
    if force or _newer(HOSTEMU_SO, [src] + _headers()):
        _run(['g++', '-O1', '-std=c++17', '-shared', '-fPIC', '-I', CSRC, '-x', 'c++', src, '-o', HOSTEMU_SO])
    return HOSTEMU_SO


HOST_TEST = os.path.join(ROOT, 'host', 'test_host')


def build_host_test(force=False):
    """host/test_host: the C++ host mirror's reference-style tests, linked against libdrynx_gpu.so."""
    src = os.path.join(ROOT, 'host', 'test_host.cpp')
    deps = [src, os.path.join(ROOT, 'host', 'drynx_host.hpp'), os.path.join(ROOT, 'include', 'drynx_gpu.h')]
    build_gpu()
    if force or _newer(HOST_TEST, deps + [GPU_SO]):
        libdir = os.path.dirname(GPU_SO)
        _run(['g++', '-O2', '-ffp-contract=off', '-std=c++17', src, '-o', HOST_TEST, '-L', libdir, '-ldrynx_gpu', '-Wl,-rpath,' + libdir,
              '-Wl,-rpath,$ORIGIN/../drynx_b200', '-L/usr/local/cuda/lib64', '-lcudart', '-Wl,-rpath,/usr/local/cuda/lib64'])
    return HOST_TEST


def build_all(force=False):
    with ThreadPoolExecutor(max_workers=3) as ex:
        futs = [ex.submit(build_host_test, force), ex.submit(build_hostemu, force)]
        if os.path.exists(os.path.join(ROOT, 'oracle', 'oracle.cpp')):
            futs.append(ex.submit(build_oracle, force))
        return [f.result() for f in futs]


if __name__ == '__main__':
    what = sys.argv[1] if len(sys.argv) > 1 else 'all'
    force = '--force' in sys.argv
    print({'gpu': build_gpu, 'oracle': build_oracle, 'hostemu': build_hostemu, 'host': build_host_test, 'all': build_all}[what](force))
