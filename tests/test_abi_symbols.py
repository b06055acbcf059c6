"""The C-ABI library loads on a CPU-only box and exports every symbol include/th_b200.h declares."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _declared():
    text = open(os.path.join(ROOT, 'include', 'th_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(th_[a-z0-9_]+)\s*\(', text)))


def test_header_symbols_are_exported(lib_built):
    lib = ctypes.CDLL(lib_built)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), 'missing export %s' % n


def test_binding_covers_header(lib_built):
    from tensorhive_fixed_b200 import native
    assert sorted(native.SIGNATURES) == _declared()
    lib = native.load()
    assert lib.th_version() == 1


def test_library_is_sm100a_only(lib_built):
    import subprocess
    out = subprocess.run(['cuobjdump', '-lelf', lib_built], capture_output=True, text=True).stdout
    archs = set(re.findall(r'sm_(\d+a?)', out))
    assert archs == {'100a'}, archs


def test_argument_errors_do_not_need_a_gpu(lib_built):
    from tensorhive_fixed_b200 import native
    lib = native.load()
    rc = lib.th_snapshot_info(None, None, None, None, None)
    assert rc == -1
    assert b'NULL' in lib.th_last_error()


def test_header_is_plain_c(tmp_path):
    """include/th_b200.h must compile as C99 (no C++ or torch types in the signatures) and link against the library."""
    import subprocess
    src = tmp_path / 'use.c'
    src.write_text('#include "th_b200.h"\n#include <stdio.h>\n'
                   'int main(void) { printf("%d %s\\n", th_version(), th_last_error()); return th_version() == TH_ABI_VERSION ? 0 : 1; }\n')
    subprocess.run(['gcc', '-std=c99', '-Wall', '-Wextra', '-Werror', '-fsyntax-only', '-x', 'c',
                    os.path.join(ROOT, 'include', 'th_b200.h')], check=True)
    subprocess.run(['gcc', '-std=c99', '-Wall', '-Werror', '-I', os.path.join(ROOT, 'include'), '-c', str(src), '-o',
                    str(tmp_path / 'use.o')], check=True)


def test_a_missing_library_fails_loudly_and_nothing_falls_back(tmp_path):
    """No CPU path hides behind the product API: with the library absent every entry raises ImportError (and the package
    itself never imports oracle/)."""
    import subprocess
    import sys
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "import tensorhive_fixed_b200 as th\n"
        "assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules)\n"
        "z = np.zeros(1, np.int64)\n"
        "for call in (lambda: th.Snapshot(0), lambda: th.usage_reduce(np.array([0, 1]), z, z, z, z, z, z)):\n"
        "    try:\n"
        "        call()\n"
        "    except ImportError as e:\n"
        "        assert 'no CPU' in str(e) or 'missing' in str(e)\n"
        "    else:\n"
        "        raise SystemExit('call succeeded without the library')\n"
        "print('loud')\n"
    ) % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TH_B200_LIB=str(tmp_path / 'nowhere' / 'libth_b200.so'))
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0 and 'loud' in out.stdout, out.stdout + out.stderr
