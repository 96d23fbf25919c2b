// selftest.cpp -- compiles and runs code shaped like the reference emitter's output against the drop-in
// headers (aquery2_b200/host/server/*.h) and libaqgpu.so.  Three post-processors:
//   dll_unchanged   the per-group loop exactly as engine/ast.py:334-460, 620-793 emits it
//                   (HashTableFactory::get, g.keys / g.values / g.offsets, col[i] = sum(v[val]), 3-argument avgw)
//   dll_legacy      the previous generation (mem_opt.cpp:31-63): AQHashTable + hashtable_push + ht_postproc
//   dll_whole       the whole-column form the GPU variant of the emitter produces
// The expected values are computed here with plain loops over the reference semantics (first-appearance
// group order, descending row order inside a group).  Needs a CUDA device: exits 2 without one.
#include "./server/libaquery.h"
#include "./server/hasher.h"
#include "./server/monetdb_conn.h"
#include "./server/aggregations.h"
#include <algorithm>
#include <cassert>
#include <map>

struct VecSource : DataSource {                      // a data source in the reference's shape (DataSource_conn.h:27-54)
    std::vector<void*> cols;
    void connect(Context* c) override { cxt = c; status = true; }
    void exec(const char*) override {}
    void* getCol(int i, int) override { return cols[i]; }
    void getDSTable(const char*, void*) override {}
    void close() override { status = false; }
    bool haserror() override { return last_error != nullptr; }
};

static int fails = 0;
#define EXPECT(c) do { if (!(c)) { printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #c); ++fails; } } while (0)

static TableInfo<int, __int128_t, double, int, vector_type<double>>* g_out1 = nullptr;

// ---- (1) unchanged emitter output -----------------------------------------------------------------
__AQEXPORT__(int) dll_unchanged(Context* cxt) {
    using namespace std;
    using namespace types;
    auto server = static_cast<DataSource*>(cxt->curr_server);
    auto len = server->cnt;
    auto id1_a = ColRef<int>(len, server->getCol(0, AINT32));
    auto v1_b = ColRef<int>(len, server->getCol(1, AINT32));
    auto v3_c = ColRef<float>(len, server->getCol(2, AFLOAT));
    const char* names[] = {"id1", "sumv1", "avgv3", "maxv1", "avgs3v1"};
    auto out = new TableInfo<int, __int128_t, double, int, vector_type<double>>("out_x", names);
    decltype(auto) col0 = out->get_col<0>();
    decltype(auto) col1 = out->get_col<1>();
    decltype(auto) col2 = out->get_col<2>();
    decltype(auto) col3 = out->get_col<3>();
    decltype(auto) col4 = out->get_col<4>();
    uint32_t total = id1_a.size;
    typedef record<decays<decltype(id1_a)>::value_t> rec_t;
    auto g = HashTableFactory<rec_t, transTypes<rec_t, hasher>>::get<decays<decltype(id1_a)>>(id1_a);
    auto sz = g.size;
    auto vecs = g.values;
    col0.resize(sz); col1.resize(sz); col2.resize(sz); col3.resize(sz); col4.resize(sz);
    auto buf_col4 = static_cast<double*>(calloc(total, sizeof(double)));
    for (uint32_t i = 0; i < sz; ++i) col4[i].init_from(vecs[i].size, buf_col4 + g.offsets[i]);
    GC::scratch_space = GC::gc_handle ? &(GC::gc_handle->scratch) : nullptr;
    for (uint32_t i = 0; i < sz; ++i) {
        auto& key = (*g.keys)[i];
        auto& val = vecs[i];
        col0[i] = (get<0>(key));
        col1[i] = (sum(v1_b[val]));
        col2[i] = (avg(v3_c[val]));
        col3[i] = (max(v1_b[val]));
        avgw(3, v1_b[val], col4[i]);
        if (GC::scratch_space) GC::scratch_space->release();
    }
    GC::scratch_space = nullptr;
    g_out1 = out;
    return 0;
}

// ---- (2) previous generation: AQHashTable -----------------------------------------------------------
static std::vector<int> g_legacy_keys;
static std::vector<long long> g_legacy_sums;
static std::vector<uint32_t> g_legacy_cnt;
__AQEXPORT__(int) dll_legacy(Context* cxt) {
    using namespace std;
    auto server = static_cast<DataSource*>(cxt->curr_server);
    auto len = server->cnt;
    auto id1_a = ColRef<int>(len, server->getCol(0, types::AINT32));
    auto id2_a = ColRef<int>(len, server->getCol(3, types::AINT32));
    auto v1_b = ColRef<int>(len, server->getCol(1, types::AINT32));
    uint32_t total = id1_a.size;
    typedef record<decays<decltype(id1_a)>::value_t, decays<decltype(id2_a)>::value_t> rec_t;
    AQHashTable<rec_t, transTypes<rec_t, hasher>> g{total};
    for (uint32_t i = 0; i < total; ++i) g.hashtable_push(forward_as_tuple(id1_a[i], id2_a[i]), i);
    auto vecs = g.ht_postproc(total);
    for (uint32_t i = 0; i < g.size(); ++i) {
        auto& key = g.values()[i];
        auto& val = vecs[i];
        g_legacy_keys.push_back(get<0>(key) * 1000 + get<1>(key));
        g_legacy_sums.push_back((long long)sum(v1_b[val]));
        g_legacy_cnt.push_back(val.size);
    }
    return 0;
}

// ---- (3) whole-column code, device-resident columns ---------------------------------------------------
static std::vector<int> g_w_keys, g_w_min;
static std::vector<__int128> g_w_sum;
static std::vector<double> g_w_avgs;
static std::vector<uint32_t> g_w_off;
__AQEXPORT__(int) dll_whole(Context* cxt) {
    auto server = static_cast<DataSource*>(cxt->curr_server);
    auto len = server->cnt;
    auto id1 = ColRef<int>(len, server->getCol(0, types::AINT32)).to_device();      // the HBM column store
    auto v1 = ColRef<int>(len, server->getCol(1, types::AINT32)).to_device();
    id1.populate_stats();
    auto g = aq::group_by(id1);
    auto k0 = g.key<0>();
    auto s1 = sum(v1[g]);
    auto m1 = min(v1[g]);
    auto a3 = avgs(3, v1[g]);
    g_w_keys = k0.to_host(); g_w_sum = s1.to_host(); g_w_min = m1.to_host(); g_w_avgs = a3.flat.to_host();
    g_w_off = g.offsets().to_host();
    EXPECT(s1.on_device() && a3.flat.on_device());
    return 0;
}

// ---- (4) the cached-table route (engine/ast.py:344-365): columns come from cxt->tables, HBM-resident ------
static std::vector<int> g_c_keys;
static std::vector<__int128> g_c_sum;
static std::vector<double> g_c_avg;
__AQEXPORT__(int) dll_cached(Context* cxt) {
    TableInfo<int, int, float, int>* tbl_source = static_cast<TableInfo<int, int, float, int>*>(cxt->tables["source"]);
    decltype(auto) id1 = tbl_source->get_col<0>();
    decltype(auto) v1 = tbl_source->get_col<1>();
    decltype(auto) v3 = tbl_source->get_col<2>();
    EXPECT(id1.on_device() && v1.on_device() && v3.on_device());
    auto g = aq::group_by(id1);
    auto s1 = sum(v1[g]);
    auto a3 = avg(v3[g]);
    EXPECT(s1.on_device() && a3.on_device());
    g_c_keys = g.key<0>().to_host(); g_c_sum = s1.to_host(); g_c_avg = a3.to_host();
    return 0;
}

// aggregates of a single value (server/aggregations.h:498-527)
static_assert(max(3) == 3 && min(2.5) == 2.5 && sum(7ll) == 7 && count(9) == 1 && var(4.f) == 0 && sumw(3u, 5) == 5 && ratiow(2u, 8.0) == 1 && deltas(6) == 0,
              "scalar overloads");

int main() {
    if (aqgpu_init(0) != 0) { printf("no CUDA device: %s\n", aqgpu_last_error()); return 2; }
    const uint32_t n = 100003;
    std::vector<int> id1(n), id2(n), v1(n);
    std::vector<float> v3(n);
    uint64_t x = 88172645463325252ull;
    auto rnd = [&] { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return x; };
    for (uint32_t i = 0; i < n; ++i) { id1[i] = 1 + rnd() % 100; id2[i] = 1 + rnd() % 7; v1[i] = (int)(rnd() % 11) - 3; v3[i] = (float)(rnd() % 100000) / 1000.f; }
    VecSource src;
    src.cnt = n;
    src.cols = { id1.data(), v1.data(), v3.data(), id2.data() };
    Context cxt;
    cxt.curr_server = &src;

    // expected: first-appearance order, rows of a group in descending order
    std::vector<int> order; std::map<int, std::vector<uint32_t>> rows;
    for (uint32_t i = 0; i < n; ++i) { if (!rows.count(id1[i])) order.push_back(id1[i]); rows[id1[i]].push_back(i); }

    EXPECT(dll_unchanged(&cxt) == 0);
    auto& o = *g_out1;
    EXPECT(o.get_col<0>().size == order.size());
    for (size_t gi = 0; gi < order.size() && fails < 5; ++gi) {
        const auto& r = rows[order[gi]];
        long long s = 0; double fs = 0; int mx = -1000;
        std::vector<int> desc(r.rbegin(), r.rend());
        for (uint32_t i : desc) { s += v1[i]; fs += (double)v3[i]; mx = std::max(mx, v1[i]); }
        EXPECT(o.get_col<0>()[gi] == order[gi]);
        EXPECT((long long)o.get_col<1>()[gi] == s);
        EXPECT(std::fabs(o.get_col<2>()[gi] - fs / r.size()) <= 1e-12 * std::fabs(fs / r.size()));
        EXPECT(o.get_col<3>()[gi] == mx);
        auto& cell = o.get_col<4>()[gi];
        EXPECT(cell.size == r.size());
        for (uint32_t j = 0; j < cell.size && fails < 5; ++j) {          // avgw(3, .) over the descending row list
            double w = 0; uint32_t c = 0;
            for (uint32_t t = (j >= 2 ? j - 2 : 0); t <= j; ++t) { w += v1[desc[t]]; ++c; }
            EXPECT(std::fabs(cell[j] - w / c) <= 1e-12 * (std::fabs(w / c) + 10));
        }
    }
    o.print(",", "\n", 3);

    EXPECT(dll_legacy(&cxt) == 0);
    {
        std::vector<int> ord2; std::map<int, std::pair<long long, uint32_t>> acc;
        for (uint32_t i = 0; i < n; ++i) { const int k = id1[i] * 1000 + id2[i]; if (!acc.count(k)) ord2.push_back(k); acc[k].first += v1[i]; acc[k].second++; }
        EXPECT(g_legacy_keys == ord2);
        for (size_t i = 0; i < ord2.size() && fails < 5; ++i) { EXPECT(g_legacy_sums[i] == acc[ord2[i]].first); EXPECT(g_legacy_cnt[i] == acc[ord2[i]].second); }
    }

    EXPECT(dll_whole(&cxt) == 0);
    EXPECT(g_w_keys == order);
    for (size_t gi = 0; gi < order.size() && fails < 5; ++gi) {
        const auto& r = rows[order[gi]];
        long long s = 0; int mn = 1000;
        for (uint32_t i : r) { s += v1[i]; mn = std::min(mn, v1[i]); }
        EXPECT((long long)g_w_sum[gi] == s);
        EXPECT(g_w_min[gi] == mn);
        EXPECT(g_w_off[gi + 1] - g_w_off[gi] == r.size());
        EXPECT(std::fabs(g_w_avgs[g_w_off[gi]] - (double)v1[r.back()]) < 1e-12);       // first cell = last row of the group
    }
    {   // the 'C' cache message of the GPU build: the table is ingested once, the post-processor reads cxt->tables
        const char* names[] = {"id1", "v1", "v3", "id2"};
        TableInfo<int, int, float, int> host("source", names);
        host.get_col<0>().init_from(n, id1.data()); host.get_col<1>().init_from(n, v1.data());
        host.get_col<2>().init_from(n, v3.data()); host.get_col<3>().init_from(n, id2.data());
        aq::cache_table(&cxt, "source", host);
        EXPECT(dll_cached(&cxt) == 0);
        EXPECT(g_c_keys == order);
        for (size_t gi = 0; gi < order.size() && fails < 5; ++gi) {
            const auto& r = rows[order[gi]];
            long long s = 0; double fs = 0;
            for (uint32_t i : r) { s += v1[i]; fs += (double)v3[i]; }
            EXPECT((long long)g_c_sum[gi] == s);
            EXPECT(std::fabs(g_c_avg[gi] - fs / r.size()) <= 1e-12 * std::fabs(fs / r.size()));
        }
    }
    {   // non-grouped forms (tests/moving_avg.a:7 `avgs(3, sales)`): the whole column is one group in table order
        ColRef<int> sales((uint32_t)n, v1.data());
        auto ma = avgs(3, sales);                       // ColRef<double>, one cell per row
        auto rs = sums(sales);
        auto dl = deltas(sales);
        EXPECT(ma.size == n && rs.size == n && dl.size == n);
        long long run = 0;
        for (uint32_t i = 0; i < n && fails < 5; ++i) {
            run += v1[i];
            const uint32_t c = i < 3 ? i + 1 : 3;
            double w = 0; for (uint32_t j = i + 1 - c; j <= i; ++j) w += v1[j];
            EXPECT(std::fabs(ma[i] - w / c) <= 1e-9);
            EXPECT((long long)rs[i] == run);
            EXPECT(dl[i] == (i ? v1[i] - v1[i - 1] : 0));
        }
        EXPECT((long long)sum(sales) == run);
        int mx = v1[0]; for (uint32_t i = 1; i < n; ++i) mx = std::max(mx, v1[i]);
        EXPECT(max(sales) == mx);
    }
    {   // string / time / timestamp key columns (benchmark/quries/Aquery/load_data.a:1 `stocksymbol STRING`; hasher.h:97-132)
        const char* syms[] = {"IBM", "AAPL", "", "A", "AA", "MSFT", "IBM.N", "GOOG"};
        std::vector<std::string> heap(n);                           // the data source's string heap: every row its own copy
        std::vector<astring_view> sym(n);
        std::vector<types::time_t> tm(n);
        std::vector<types::timestamp_t> ts(n);
        for (uint32_t i = 0; i < n; ++i) {
            heap[i] = syms[rnd() % 8]; sym[i] = astring_view(heap[i].c_str());
            tm[i] = types::time_t((unsigned)(rnd() % 3) * 500u, (unsigned char)(rnd() % 2), 0, (unsigned char)(rnd() % 3));
            reinterpret_cast<unsigned char*>(&tm[i])[7] = (unsigned char)rnd();      // padding garbage: not part of the key
            ts[i] = types::timestamp_t(types::date_t((unsigned char)(1 + rnd() % 2), 7, 2024), tm[i]);
        }
        ColRef<astring_view> c_sym(n, sym.data());
        ColRef<types::time_t> c_tm(n, tm.data());
        ColRef<types::timestamp_t> c_ts(n, ts.data());
        ColRef<int> c_v(n, v1.data());
        ColRef<int> c_id2(n, id2.data());
        {   // (sym, id2) -> sum(v1), the unchanged per-group loop
            auto g = HashTableFactory<record<astring_view, int>, transTypes<record<astring_view, int>, hasher>>::get(c_sym, c_id2);
            std::vector<std::pair<std::string, int>> ord; std::map<std::pair<std::string, int>, long long> acc;
            for (uint32_t i = 0; i < n; ++i) { auto k = std::make_pair(heap[i], id2[i]); if (!acc.count(k)) ord.push_back(k); acc[k] += v1[i]; }
            EXPECT(g.size == ord.size());
            for (uint32_t i = 0; i < g.size && fails < 5; ++i) {
                auto& key = (*g.keys)[i];
                EXPECT(std::string(std::get<0>(key).rstr) == ord[i].first && std::get<1>(key) == ord[i].second);
                EXPECT((long long)sum(c_v[g.values[i]]) == acc[ord[i]]);
            }
        }
        {   // time key, whole-column form
            auto g = aq::group_by(c_tm);
            auto img = [](const types::time_t& t) { return ((unsigned long long)t.hours << 48) | ((unsigned long long)t.minutes << 40) | ((unsigned long long)t.seconds << 32) | t.ms; };
            std::vector<unsigned long long> ord; std::map<unsigned long long, long long> acc;
            for (uint32_t i = 0; i < n; ++i) { auto k = img(tm[i]); if (!acc.count(k)) ord.push_back(k); acc[k] += v1[i]; }
            auto keys = g.key<0>(); auto s = sum(c_v[g]).to_host();
            EXPECT(g.size == ord.size() && g.size == 18);
            for (uint32_t i = 0; i < g.size && fails < 5; ++i) { EXPECT(img(keys[i]) == ord[i]); EXPECT((long long)s[i] == acc[ord[i]]); }
        }
        {   // timestamp key
            auto g = aq::group_by(c_ts);
            std::vector<uint32_t> firsts; std::vector<types::timestamp_t> seen;
            for (uint32_t i = 0; i < n; ++i) { bool f = false; for (auto& t : seen) if (t == ts[i]) { f = true; break; } if (!f) { seen.push_back(ts[i]); firsts.push_back(i); } }
            auto keys = g.key<0>();
            EXPECT(g.size == seen.size() && g.size == 36);
            for (uint32_t i = 0; i < g.size && fails < 5; ++i) EXPECT(keys[i] == seen[i]);
        }
    }
    {   // element-wise operators (table.h:826-973) and tests/network.a:8-11: GROUP BY src, sums(deltas(_time) > 120)
        std::vector<int> tm(n); std::vector<std::string> heap(n); std::vector<astring_view> srcv(n);
        const char* hosts[] = {"a1", "b2", "c3", "a1b"};
        int clock = 0;
        for (uint32_t i = 0; i < n; ++i) { clock += 1 + (int)(rnd() % 200); tm[i] = clock; heap[i] = hosts[rnd() % 4]; srcv[i] = astring_view(heap[i].c_str()); }
        ColRef<int> c_t(n, tm.data()), c_v(n, v1.data());
        ColRef<float> c_f(n, v3.data());
        ColRef<astring_view> c_src(n, srcv.data());
        auto gap = deltas(c_t) > 120;                       // ColRef<bool>
        auto sess = sums(gap);                              // ColRef<long long>: running count of gaps = session id
        auto g = aq::group_by(c_src, sess);
        auto cnt = g.counts();
        auto av = avg(c_v[g]).to_host();
        std::vector<std::pair<std::string, long long>> ord; std::map<std::pair<std::string, long long>, std::pair<long long, uint32_t>> acc;
        long long sid = 0;
        for (uint32_t i = 0; i < n; ++i) {
            if (i && tm[i] - tm[i - 1] > 120) ++sid;
            EXPECT(fails > 5 || sess[i] == sid);
            auto k = std::make_pair(heap[i], sid);
            if (!acc.count(k)) ord.push_back(k);
            acc[k].first += v1[i]; acc[k].second++;
        }
        EXPECT(g.size == ord.size());
        auto k0 = g.key<0>(); auto k1 = g.key<1>();
        for (uint32_t i = 0; i < g.size && fails < 5; ++i) {
            EXPECT(std::string(k0[i].rstr) == ord[i].first && k1[i] == ord[i].second);
            EXPECT(cnt[i] == acc[ord[i]].second);
            EXPECT(std::fabs(av[i] - (double)acc[ord[i]].first / acc[ord[i]].second) <= 1e-12 * (1 + std::fabs(av[i])));
        }
        // result element types and cells of the infix forms
        auto s1 = c_v + c_f;   static_assert(std::is_same<decltype(s1), ColRef<float>>::value, "int + float -> float");
        auto s2 = c_v * 3;     static_assert(std::is_same<decltype(s2), ColRef<__int128>>::value, "int column * scalar -> __int128 (GetLongType)");
        auto s3 = c_v / 2;     static_assert(std::is_same<decltype(s3), ColRef<float>>::value, "int column / scalar -> float (GetFPType), integer division inside");
        auto s4 = 10 - c_v;    static_assert(std::is_same<decltype(s4), ColRef<int>>::value, "scalar - int column -> int");
        auto s5 = c_v * c_v;   static_assert(std::is_same<decltype(s5), ColRef<int>>::value, "column * column -> Coercion (vector_type.hpp:386-396)");
        auto s6 = c_f >= 50.f; static_assert(std::is_same<decltype(s6), ColRef<bool>>::value, "comparison -> bool");
        auto sq = sqrt(c_f); auto tr = truncate(c_f, 1);
        static_assert(std::is_same<decltype(sq), ColRef<double>>::value && std::is_same<decltype(tr), ColRef<float>>::value, "sqrt -> double, truncate -> T");
        for (uint32_t i = 0; i < n && fails < 5; i += 13) {
            EXPECT(sq[i] == (double)std::sqrt(v3[i]));
            EXPECT(tr[i] == (float)(std::round((double)v3[i] * 10.0) / 10.0));
        }
        ColRef<signed char> r8(n); aqop_lt(c_v, c_t, r8);   // root of an emitted expression: aqop_lt(lhs, rhs, ret) with the emitter's byte type
        for (uint32_t i = 0; i < n && fails < 5; ++i) {
            EXPECT(s1[i] == (float)v1[i] + v3[i]);
            EXPECT((long long)s2[i] == (long long)v1[i] * 3);
            EXPECT(s3[i] == (float)(v1[i] / 2));
            EXPECT(s4[i] == 10 - v1[i]);
            EXPECT(s5[i] == v1[i] * v1[i]);
            EXPECT(s6[i] == (v3[i] >= 50.f));
            EXPECT(r8[i] == (v1[i] < tm[i] ? 1 : 0));
        }
    }
    {   // TableInfo::order_by / materialize / distinct (table.h:429-461, 601-614) and the gather col[idx] (table.h:184-189)
        const char* names[] = {"id1", "id2", "v3"};
        TableInfo<int, int, float> t("t", names);
        t.get_col<0>().init_from(n, id1.data()); t.get_col<1>().init_from(n, id2.data()); t.get_col<2>().init_from(n, v3.data());
        auto* ord = t.order_by<0, -2>();                                   // ASC id1, DESC id2; ties keep table order
        std::vector<uint32_t> want(n);
        for (uint32_t i = 0; i < n; ++i) want[i] = i;
        std::stable_sort(want.begin(), want.end(), [&](uint32_t a, uint32_t b) { return std::make_tuple(id1[a], -id2[a]) < std::make_tuple(id1[b], -id2[b]); });
        EXPECT(ord->size == n);
        for (uint32_t i = 0; i < n && fails < 5; ++i) EXPECT((*ord)[i] == want[i]);
        vector_type<uint32_t> part(1000);                                  // a row list (one group's rows): sorted in place
        for (uint32_t i = 0; i < 1000; ++i) part[i] = (uint32_t)((i * 7919u) % n);
        std::vector<uint32_t> pw(part.container, part.container + 1000);
        std::stable_sort(pw.begin(), pw.end(), [&](uint32_t a, uint32_t b) { return v3[a] > v3[b]; });
        t.order_by<-3>(&part);                                             // DESC v3
        for (uint32_t i = 0; i < 1000 && fails < 5; ++i) EXPECT(part[i] == pw[i]);
        auto* sorted = t.materialize_copy(*ord);
        EXPECT(sorted->get_col<0>().size == n);
        for (uint32_t i = 0; i < n && fails < 5; ++i) { EXPECT(sorted->get_col<0>()[i] == id1[want[i]]); EXPECT(sorted->get_col<2>()[i] == v3[want[i]]); }
        TableInfo<int, int> d("d", names);
        d.get_col<0>().init_from(n, id1.data()); d.get_col<1>().init_from(n, id2.data());
        d.distinct();
        std::vector<std::pair<int, int>> seen; std::map<std::pair<int, int>, int> have;
        for (uint32_t i = 0; i < n; ++i) { auto k = std::make_pair(id1[i], id2[i]); if (!have.count(k)) { have[k] = 1; seen.push_back(k); } }
        EXPECT(d.get_col<0>().size == seen.size() && d.get_col<1>().size == seen.size());
        for (size_t i = 0; i < seen.size() && fails < 5; ++i) EXPECT(d.get_col<0>()[i] == seen[i].first && d.get_col<1>()[i] == seen[i].second);
        delete ord;
    }
    {   // N4: an elemental-return UDF in the emitter's shape (engine/ast.py:768-777; tests/udf3.a) applied per group, and a
        // module function taking ColRef by value (sdk/example.cpp:4) fed from the HBM store
        auto running_cov = [](auto x, auto y, auto win, auto _builtin_len, auto& _builtin_ret) {     // generic lambda as emitted
            double xs = 0, ys = 0, xy = 0;
            for (uint32_t i = 0; i < _builtin_len; ++i) {
                xs += x[i]; ys += y[i]; xy += (double)x[i] * y[i];
                if (i >= (uint32_t)win) { xs -= x[i - win]; ys -= y[i - win]; xy -= (double)x[i - win] * y[i - win]; }
                const double m = i < (uint32_t)win ? i + 1 : win;
                _builtin_ret[i] = xy / m - (xs / m) * (ys / m);
            }
            return 0;
        };
        ColRef<int> c_k(n, id2.data()), c_a(n, v1.data());
        ColRef<float> c_b(n, v3.data());
        auto g = HashTableFactory<record<int>, transTypes<record<int>, hasher>>::get(c_k);
        std::vector<std::vector<double>> outs(g.size);
        for (uint32_t i = 0; i < g.size; ++i) {
            auto& val = g.values[i];
            const uint32_t len = val.size;                       // _builtin_len (define_len_var)
            outs[i].resize(len);
            running_cov(c_a[val], c_b[val], 4, len, outs[i]);
        }
        std::map<int, std::vector<uint32_t>> rows2; std::vector<int> ord2;
        for (uint32_t i = 0; i < n; ++i) { if (!rows2.count(id2[i])) ord2.push_back(id2[i]); rows2[id2[i]].push_back(i); }
        EXPECT(g.size == ord2.size());
        for (uint32_t gi = 0; gi < g.size && fails < 5; ++gi) {
            std::vector<uint32_t> desc(rows2[ord2[gi]].rbegin(), rows2[ord2[gi]].rend());     // descending row order inside a group
            EXPECT(outs[gi].size() == desc.size());
            std::vector<int> xa; std::vector<float> ya;
            for (uint32_t r : desc) { xa.push_back(v1[r]); ya.push_back(v3[r]); }
            std::vector<double> want(desc.size());
            running_cov(xa, ya, 4, (uint32_t)desc.size(), want);
            for (size_t j = 0; j < desc.size() && fails < 5; j += 97) EXPECT(outs[gi][j] == want[j]);
        }
        auto dev = c_b.to_device();
        auto mulvec = [](int a, ColRef<float> b) { double s = 0; for (uint32_t i = 0; i < b.size; ++i) s += a * b.container[i]; return s; };   // CPU dereference
        double want = 0; for (uint32_t i = 0; i < n; ++i) want += 2 * v3[i];
        EXPECT(dev.on_device());
        EXPECT(mulvec(2, aq::host_view(dev)) == want);
        EXPECT(mulvec(2, aq::host_view(c_b)) == want);
    }
    printf(fails ? "selftest: %d FAILED\n" : "selftest ok (%d failures)\n", fails);
    return fails ? 1 : 0;
}
