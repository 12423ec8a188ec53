"""The C ABI is one surface seen from four places -- the header, the exported symbols of the two libraries, the ctypes
bindings and the documentation.  They drift apart silently (a function declared but never exported fails only at the
caller's link; one missing from ``_lib.py`` is untestable from Python), so this pins them to each other.  CPU only:
both libraries are cross-compiled here.  The reference has no API at all (two ``main()`` programs, MDF_kernel.cu:101,
kernel.cu:148); this is the surface a user of it switches to."""
import os
import re
import subprocess


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mcp_b200.h")
# declared for plain-C hosts that do not link the CUDA runtime (examples/solve_mdf.c); Python uses torch instead
C_HOST_ONLY = {"mcp_set_device", "mcp_device_synchronize", "mcp_copy_to_host", "mcp_copy_to_device"}


def _declared():
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return set(re.findall(r"\b(mcp_[a-z0-9_]+)\s*\(", text))


def _exported():
    from mpi_cuda_process_b200.ops import build
    out = set()
    for path in (build.build_cuda(), build.build_host()):
        nm = subprocess.run(["nm", "-D", "--defined-only", path], capture_output=True, text=True, check=True).stdout
        out |= set(re.findall(r" T (mcp_[a-z0-9_]+)", nm))
    return out


def test_header_and_libraries_agree():
    decl, exp = _declared(), _exported()
    assert decl - exp == set(), f"declared in mcp_b200.h but not exported: {sorted(decl - exp)}"
    assert exp - decl == set(), f"exported but not declared in mcp_b200.h: {sorted(exp - decl)}"
    assert len(decl) >= 70


def test_every_entry_point_has_a_ctypes_binding():
    src = open(os.path.join(ROOT, "mpi_cuda_process_b200", "ops", "_lib.py")).read()
    missing = sorted(f for f in _declared() - C_HOST_ONLY if not re.search(r"\b" + f + r"\b", src))
    assert missing == []


def test_every_entry_point_is_documented():
    docs = "".join(open(os.path.join(ROOT, p)).read() for p in ("docs/API.md", "INTEGRATION.md"))
    missing = sorted(f for f in _declared() if not re.search(r"\b" + f + r"\b", docs))
    assert missing == []


def test_c_example_uses_only_declared_functions():
    src = open(os.path.join(ROOT, "examples", "solve_mdf.c")).read()
    used = set(re.findall(r"\b(mcp_[a-z0-9_]+)\s*\(", src))
    assert used and used <= _declared()
    assert len(C_HOST_ONLY & used) >= 3      # the reason those exist


def test_header_is_strict_c99_and_cxx11_and_defines_the_return_codes(tmp_path):
    """A host in plain C must be able to include the header on its own and name every return code."""
    body = ('#include "mcp_b200.h"\n'
            "int main(void) { int codes[] = {MCP_OK, MCP_ERR_CUDA, MCP_ERR_ARG, MCP_ERR_UNSUPPORTED, MCP_ERR_TIMEOUT,\n"
            "                                MCP_ERR_STATE, MCP_STOP_SLOT_LIMIT, MCP_OPT_TIMEOUT_MS, MCP_RES_FINAL};\n"
            "  return (int)(sizeof(codes) / sizeof(codes[0])) == 9 && sizeof(mcp_fused2d_desc) > 0 ? 0 : 1; }\n")
    inc = os.path.join(ROOT, "include")
    for name, cc, std in (("h.c", "gcc", "-std=c99"), ("h.cpp", "g++", "-std=c++11")):
        src = tmp_path / name
        src.write_text(body)
        r = subprocess.run([cc, std, "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", inc, "-c", str(src),
                            "-o", str(tmp_path / (name + ".o"))], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
