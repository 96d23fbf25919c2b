// harness.cpp -- stands in for the reference's server.so around ONE generated translation unit: builds a Context
// and a DataSource whose columns come from col_<i>.bin files, calls the generated c(cxt) (server/server.cpp:296-307).
// The translation unit is #included so that one g++ call builds the program; tests/test_codegen.py drives it.
#include QUERY_TU
#include <cstdio>
#include <vector>
#include <string>

struct FileSource : DataSource {
    std::vector<std::vector<char>> cols;
    void connect(Context* c) override { cxt = c; status = true; }
    void exec(const char*) override {}
    void* getCol(int i, int) override { return cols[i].data(); }
    void getDSTable(const char*, void*) override {}
    void close() override {}
    bool haserror() override { return false; }
};

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    const long long n = atoll(argv[1]);
    const int ncols = atoi(argv[2]);
    if (aqgpu_init(0) != 0) { printf("no CUDA device: %s\n", aqgpu_last_error()); return 3; }
    FileSource src;
    src.cnt = n;
    for (int i = 0; i < ncols; ++i) {
        std::vector<char> buf(size_t(n) * 4);
        FILE* f = fopen(("col_" + std::to_string(i) + ".bin").c_str(), "rb");
        if (f) { if (fread(buf.data(), 4, n, f) != (size_t)n) return 4; fclose(f); }
        src.cols.push_back(std::move(buf));
    }
    Context cxt;
    cxt.curr_server = &src;
    return QUERY_ENTRY(&cxt);
}
