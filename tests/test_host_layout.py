"""The drop-in Context / DataSource (aquery2_b200/host/aqgpu_host.hpp) against the reference's own headers:
member offsets, sizes and virtual-function slots must be identical, or a dll.so built on the drop-in headers
reads garbage when the reference's server.so calls c(cxt) (server/server.cpp:296-307).

Runs where /root/reference is mounted (this container); the offsets are also pinned by static_asserts inside
the header, which every build of the host side checks.  Second test: an emitter-shaped translation unit that
lies NEXT TO a server/ directory (quoted includes search the includer's directory first) compiles against the
drop-in headers with the documented -I recipe."""
import os
import shutil
import subprocess
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
HOST = os.path.join(ROOT, "aquery2_b200", "host")

PROBE = r"""
#include <cstdio>
#include <cstddef>
#pragma GCC diagnostic ignored "-Winvalid-offsetof"
struct Probe : DataSource {
    void connect(Context*) override {}
    void exec(const char*) override {}
    void* getCol(int, int) override { return (void*)0x1234; }
    void getDSTable(const char*, void*) override {}
    void close() override {}
    bool haserror() override { return true; }
};
int main() {
#define O(T, m) printf(#T "." #m " %zu\n", offsetof(T, m))
    O(DataSource, server); O(DataSource, cxt); O(DataSource, status); O(DataSource, query); O(DataSource, DataSourceType);
    O(DataSource, res); O(DataSource, ret_col); O(DataSource, cnt); O(DataSource, last_error); O(DataSource, handle);
    printf("sizeof(DataSource) %zu\n", sizeof(DataSource));
    O(Context, module_function_maps); O(Context, cfg); O(Context, n_buffers); O(Context, sz_bufs); O(Context, buffers);
    O(Context, curr_server); O(Context, alt_server); O(Context, log_level); O(Context, current); O(Context, aquery_root_path);
    O(Context, thread_pool); O(Context, gc); O(Context, print); O(Context, tables); O(Context, cols); O(Context, stored_proc);
    O(Context, triggers); O(Context, it_host); O(Context, ct_host);
    printf("sizeof(Context) %zu\n", sizeof(Context));
    printf("sizeof(Session) %zu sizeof(StoredProcedure) %zu sizeof(Config) %zu\n", sizeof(Session), sizeof(StoredProcedure), sizeof(Config));
    // virtual slots: call through the base pointer by slot number, the way foreign code compiled on the OTHER header would
    Probe p; DataSource* d = &p;
    void** vt = *reinterpret_cast<void***>(d);
    typedef void* (*getcol_t)(DataSource*, int, int);
    typedef bool (*haserr_t)(DataSource*);
    printf("slot2 getCol %p slot5 haserror %d\n", reinterpret_cast<getcol_t>(vt[2])(d, 0, 0), (int)reinterpret_cast<haserr_t>(vt[5])(d));
    return 0;
}
"""


def _run(tmp_path, name, header_lines, flags):
    src = tmp_path / f"{name}.cpp"
    src.write_text("\n".join(header_lines) + PROBE)
    exe = tmp_path / name
    r = subprocess.run(["g++", "-std=c++20", "-O0", "-w", *flags, str(src), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    return out.stdout


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "server")), reason="reference sources not mounted")
def test_context_and_datasource_match_the_reference_layout(tmp_path):
    ours = _run(tmp_path, "ours", ['#include "server/libaquery.h"', '#include "server/monetdb_conn.h"'], [f"-I{HOST}"])
    # the reference's default build: THREADING=1 (aquery_config.py:29)
    theirs = _run(tmp_path, "theirs", ['#include "server/libaquery.h"', '#include "server/DataSource_conn.h"'], ["-DTHREADING", f"-I{REF}"])
    assert ours == theirs, "\n--- drop-in ---\n" + ours + "\n--- reference ---\n" + theirs


def test_emitter_shaped_tu_next_to_a_server_directory(tmp_path):
    """out.cpp as engine/ast.py writes it ('#include "./server/libaquery.h"', header.cxx:1, engine/ast.py:244,629)
    in a directory that HAS a server/ sub-directory: the recipe of INTEGRATION.md copies the drop-in headers there."""
    work = tmp_path / "aquery_root"
    shutil.copytree(os.path.join(HOST, "server"), work / "server")
    shutil.copy(os.path.join(HOST, "aqgpu_host.hpp"), work / "aqgpu_host.hpp")
    (work / "out.cpp").write_text(textwrap.dedent('''
        #include "./server/libaquery.h"
        #include "./server/monetdb_conn.h"
        #include "./server/aggregations.h"
        #include "./server/hasher.h"
        extern "C" int dll_probe(Context* cxt) {
            auto server = static_cast<DataSource*>(cxt->curr_server);
            auto len = server->cnt;
            auto col = ColRef<int>(len, server->getCol(0, 0));
            (void)col;
            return 0;
        }
    '''))
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-w", "-fPIC", "-shared", f"-I{os.path.join(ROOT, 'include')}", "out.cpp", "-o", "dll.so"],
                       cwd=work, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
