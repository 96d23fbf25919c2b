// aqgpu_host.hpp -- drop-in host side of AQuery++'s post-processor API over libaqgpu (include/aqgpu.h).
//
// The generated post-processor (reference engine/ast.py:334-460, 620-793; sample mem_opt.cpp:16-71)
// and SDK modules compile against the reference's headers server/{vector_type.hpp, table.h, hasher.h,
// aggregations.h, types.h, gc.h, libaquery.h}.  The files of the same names next to this one forward
// here; this header keeps the reference's identifiers and argument meaning:
//
//   vector_type<T>, ColRef<T>, TableInfo<Ts...>        server/vector_type.hpp:30-431, table.h:68-258, 380-617
//   record<>, decays<>, transTypes<,hasher>, value_type<>   server/types.h:356-450, hasher.h:66-140
//   HashTableFactory<K,H>::get(cols...) -> {size, keys, values, offsets}   server/hasher.h:201-207, 324-357
//   AQHashTable<K,H>: hashtable_push / hashtable_push_all / ht_postproc    server/hasher.h:146-199
//   count sum avg min max first last, sums avgs mins maxs, sumw avgw minw maxw (SQL sums(w,x) ...),
//   deltas prev aggnext ratios ratiow, var stddev corr, vars stddevs varw stddevw     server/aggregations.h
//
// A column's `container` may point to host memory or into the HBM column store (ColRef::to_device());
// every library call accepts both.  Two ways to use it:
//   (1) unchanged generated code: the per-group loop `col[i] = sum(v[vecs[i]])` keeps compiling; `vecs[i]`
//       is a proxy of group i and the FIRST call of an aggregate over a column evaluates it for ALL
//       groups on the GPU (one aqgpu_group_reduce / aqgpu_group_scan) and memoises the result column, so
//       the loop degenerates into reads of that column;
//   (2) whole-column code (what the GPU variant of the emitter produces):
//       `auto g = aq::group_by(id1, id2); out0 = g.key<0>(); out1 = sum(v1[g]); out2 = avgs(5, price[g]);`
//
// Errors: a failed library call throws aq::error (the reference has no error path inside c(cxt);
// the exported function may catch and return non-zero).  There is no CPU evaluation path in here:
// without libaqgpu and a CUDA device every aggregate throws.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <memory>
#include <map>
#include <stdexcept>
#include <string>
#include <tuple>
#include <type_traits>
#include <unordered_map>
#include <utility>
#include <vector>
#include <iostream>
#if defined(__has_include) && !__has_include("../../include/aqgpu.h")
#include <aqgpu.h>                      // headers copied next to a generated out.cpp: -I <repo>/include
#else
#include "../../include/aqgpu.h"
#endif

#ifndef __AQEXPORT__
#define __AQEXPORT__(_Ty) extern "C" _Ty __attribute__((visibility("default")))      // server/libaquery.h:184
#endif

// ---- types (server/types.h, server/aquery_types.h) -------------------------------------------------
namespace types {
enum Type_t { AINT32, AFLOAT, ASTR, ADOUBLE, ALDOUBLE, AINT64, AINT128, AINT16, ADATE, ATIME, AINT8, AUINT32, AUINT64,
              AUINT128, AUINT16, AUINT8, ABOOL, VECTOR, ATIMESTAMP, ACHAR, ASV, NONE, ERROR };
template <class T> struct Types { static constexpr Type_t getType() { return ERROR; } };
#define AQ_TYPE(T, E) template <> struct Types<T> { static constexpr Type_t getType() { return E; } };
AQ_TYPE(int, AINT32) AQ_TYPE(float, AFLOAT) AQ_TYPE(double, ADOUBLE) AQ_TYPE(long long, AINT64) AQ_TYPE(long, AINT64)
AQ_TYPE(__int128, AINT128) AQ_TYPE(unsigned int, AUINT32) AQ_TYPE(unsigned long, AUINT64) AQ_TYPE(unsigned long long, AUINT64)
#undef AQ_TYPE
// server/types.h:199-210: sum(int) is __int128, sum(fp) is double; avg is double; ratios(int32) is float
template <class T> using GetLongType = std::conditional_t<std::is_floating_point<std::decay_t<T>>::value, double,
                                       std::conditional_t<std::is_unsigned<std::decay_t<T>>::value, unsigned __int128, __int128>>;
template <class T> using GetFPType = std::conditional_t<sizeof(std::decay_t<T>) == sizeof(float), float, double>;
// key-capable cell types, bit-compatible with server/types.h:82-86,105-109,128-130
struct date_t {
    unsigned char day = 0, month = 0; short year = 0;
    date_t() = default;
    date_t(unsigned char d, unsigned char m, short y) : day(d), month(m), year(y) {}
    bool operator==(const date_t& o) const { return year == o.year && month == o.month && day == o.day; }      // libaquery.cpp:112-114
    bool operator!=(const date_t& o) const { return !(*this == o); }
};
struct time_t {
    unsigned int ms = 0; unsigned char seconds = 0, minutes = 0, hours = 0;
    time_t() = default;
    time_t(unsigned int ms, unsigned char s, unsigned char m, unsigned char h) : ms(ms), seconds(s), minutes(m), hours(h) {}
    bool operator==(const time_t& o) const { return hours == o.hours && minutes == o.minutes && seconds == o.seconds && ms == o.ms; }   // :162-164
    bool operator!=(const time_t& o) const { return !(*this == o); }
};
struct timestamp_t {
    date_t date; time_t time;
    timestamp_t() = default;
    timestamp_t(const date_t& d, const time_t& t) : date(d), time(t) {}
    bool operator==(const timestamp_t& o) const { return date == o.date && time == o.time; }                    // :201-203
    bool operator!=(const timestamp_t& o) const { return !(*this == o); }
};
static_assert(sizeof(date_t) == 4 && sizeof(time_t) == 8 && sizeof(timestamp_t) == 12, "cells of server/types.h");
// printed like toString (server/libaquery.cpp:86-95, 138-149, 183-188): year-month-day, hours:minutes:seconds:ms
inline std::ostream& operator<<(std::ostream& os, const date_t& v) { return os << (int)v.year << '-' << (int)v.month << '-' << (int)v.day; }
inline std::ostream& operator<<(std::ostream& os, const time_t& v) { return os << (int)v.hours << ':' << (int)v.minutes << ':' << (int)v.seconds << ':' << v.ms; }
inline std::ostream& operator<<(std::ostream& os, const timestamp_t& v) { return os << v.date << ' ' << v.time; }
template <> struct Types<date_t> { static constexpr Type_t getType() { return ADATE; } };
template <> struct Types<time_t> { static constexpr Type_t getType() { return ATIME; } };
template <> struct Types<timestamp_t> { static constexpr Type_t getType() { return ATIMESTAMP; } };
}  // namespace types

// server/types.h:280-333: a string cell is a char pointer into the data source's heap
union astring_view {
    const unsigned char* str = nullptr; const signed char* sstr; const char* rstr; size_t ptr;
    constexpr astring_view(const char* s) noexcept : rstr(s) {}
    constexpr astring_view(const unsigned char* s) noexcept : str(s) {}
    constexpr astring_view() noexcept = default;
    bool operator==(const astring_view& r) const { return std::strcmp(rstr, r.rstr) == 0; }                     // types.h:299-309
    operator const char*() const { return rstr; }
};
template <> struct types::Types<astring_view> { static constexpr Type_t getType() { return ASTR; } };

// ---- error handling -----------------------------------------------------------------------------------
namespace aq {
struct error : std::runtime_error { using std::runtime_error::runtime_error; };
inline void check(int rc) { if (rc != 0) throw error(std::string("aqgpu: ") + aqgpu_last_error()); }
template <class T> constexpr int dtype_of() {
    using D = std::decay_t<T>;
    return std::is_same<D, int>::value ? AQ_INT32 : std::is_same<D, unsigned>::value ? AQ_UINT32
         : std::is_same<D, float>::value ? AQ_FLOAT : std::is_same<D, double>::value ? AQ_DOUBLE
         : (std::is_same<D, long long>::value || std::is_same<D, long>::value) ? AQ_INT64
         : (std::is_same<D, __int128>::value || std::is_same<D, unsigned __int128>::value) ? AQ_INT128
         : std::is_same<D, bool>::value ? AQ_BOOL : -1;
}
struct GroupsHandle;
struct GroupRows;
template <class T> struct GroupVec;
template <class T> struct GroupedCol;
}  // namespace aq

// ---- GC shim (server/gc.h:7-106): the generated code sets / releases GC::scratch_space --------------------
struct ScratchSpace { void release() {} void reset() {} };
struct GC {
    static inline ScratchSpace* scratch_space = nullptr;
    static inline GC* gc_handle = nullptr;
    ScratchSpace scratch;
};

// ---- vector_type (server/vector_type.hpp:25-431) ----------------------------------------------------------
struct vectortype_cstorage { void* container; unsigned int size, capacity; };

template <class _Ty> class vector_type {
public:
    typedef _Ty value_t;
    typedef vector_type<_Ty> Decayed_t;
    _Ty* container;
    uint32_t size, capacity;                    // capacity > 0: owns `container` (host malloc or HBM)
    constexpr vector_type() noexcept : container(nullptr), size(0), capacity(0) {}
    explicit vector_type(const uint32_t& n) : container(static_cast<_Ty*>(malloc(sizeof(_Ty) * (n ? n : 1)))), size(n), capacity(n ? n : 1) {}
    vector_type(const uint32_t n, void* data) : container(static_cast<_Ty*>(data)), size(n), capacity(0) {}
    constexpr vector_type(_Ty* data, uint32_t n) noexcept : container(data), size(n), capacity(0) {}
    vector_type(std::initializer_list<_Ty> l) : vector_type((uint32_t)l.size()) { uint32_t i = 0; for (const auto& x : l) container[i++] = x; }
    vector_type(vector_type<_Ty>& o) noexcept : container(o.container), size(o.size), capacity(0) {}         // reference semantics (:96-98)
    vector_type(const vector_type<_Ty>& o) : container(o.container), size(o.size), capacity(0) {
        if (o.capacity) { alloc_like(o); copy_from(o); }
    }
    vector_type(vector_type<_Ty>&& o) noexcept : container(o.container), size(o.size), capacity(o.capacity) { o.container = nullptr; o.size = o.capacity = 0; }
    explicit vector_type(vectortype_cstorage s) noexcept : container(static_cast<_Ty*>(s.container)), size(s.size), capacity(s.capacity) {}
    ~vector_type() { release(); }
    vector_type<_Ty>& operator=(const vector_type<_Ty>& o) {
        if (this == &o) return *this;
        release(); container = o.container; size = o.size; capacity = 0;
        if (o.capacity) { alloc_like(o); copy_from(o); }
        return *this;
    }
    vector_type<_Ty>& operator=(vector_type<_Ty>&& o) noexcept {
        if (this == &o) return *this;
        release(); container = o.container; size = o.size; capacity = o.capacity;
        o.container = nullptr; o.size = o.capacity = 0;
        return *this;
    }
    void init_from(const uint32_t n, void* data) { release(); container = static_cast<_Ty*>(data); size = n; capacity = 0; }
    bool on_device() const { return container && aqgpu_ptr_is_device(container) == 1; }
    _Ty& operator[](const uint32_t i) const { return container[i]; }            // host-resident vectors only
    _Ty* begin() const { return container; }
    _Ty* end() const { return container + size; }
    _Ty& back() { return container[size - 1]; }
    void reserve(uint32_t n) { if (n > capacity_or_size()) grow_to(n); }
    void resize(uint32_t n) {                                                   // zero-fills (server/vector_type.hpp:205-207)
        const uint32_t old = size;
        if (n > capacity_or_size() || capacity == 0) grow_to(n > 1 ? n : 1);
        if (n > old) memset((void*)(container + old), 0, sizeof(_Ty) * (n - old));
        size = n;
    }
    void emplace_back(const _Ty& v) { if (capacity == 0 || size >= capacity) grow_to(size ? size * 2 : 8); container[size++] = v; }
    void clear() { size = 0; }
    vector_type<_Ty> subvec(uint32_t start, uint32_t end) const { return vector_type<_Ty>(container + start, end - start); }   // :348-355
    vector_type<_Ty> subvec(uint32_t start = 0) const { return subvec(start, size); }
    vector_type<_Ty> subvec_deep(uint32_t start, uint32_t end) const { vector_type<_Ty> r(end - start); memcpy(r.container, container + start, sizeof(_Ty) * (end - start)); return r; }
    vector_type<_Ty> getRef() { return vector_type<_Ty>(container, size); }
    vectortype_cstorage s() const { return vectortype_cstorage{container, size, capacity}; }
    // host copy of a column that may live in HBM
    std::vector<_Ty> to_host() const {
        std::vector<_Ty> h(size);
        if (size) { if (on_device()) aq::check(aqgpu_col_download(h.data(), container, sizeof(_Ty) * size)); else memcpy(h.data(), container, sizeof(_Ty) * size); }
        return h;
    }
    void out(uint32_t n = 4000, const char* sep = " ") const {
        const auto h = to_host();
        n = n < size ? n : size;
        std::cout << '(';
        for (uint32_t i = 0; i < n; ++i) { if (i) std::cout << sep; print_cell(h[i]); }
        std::cout << (n < size ? " ... )" : ")");
    }
    static void print_cell(const _Ty& v) {
        if constexpr (std::is_same<_Ty, __int128>::value || std::is_same<_Ty, unsigned __int128>::value) std::cout << (long long)v;
        else std::cout << v;
    }
protected:
    uint32_t capacity_or_size() const { return capacity ? capacity : 0; }
    void alloc_like(const vector_type<_Ty>& o) {
        if (o.on_device()) { void* p = nullptr; aq::check(aqgpu_col_alloc(sizeof(_Ty) * (o.size ? o.size : 1), &p)); container = static_cast<_Ty*>(p); }
        else container = static_cast<_Ty*>(malloc(sizeof(_Ty) * (o.size ? o.size : 1)));
        capacity = o.size ? o.size : 1;
    }
    void copy_from(const vector_type<_Ty>& o) {
        if (!o.size) return;
        if (o.on_device()) { std::vector<_Ty> h = o.to_host(); aq::check(aqgpu_col_upload(container, h.data(), sizeof(_Ty) * o.size)); }
        else memcpy(container, o.container, sizeof(_Ty) * o.size);
    }
    void grow_to(uint32_t n) {
        _Ty* p = static_cast<_Ty*>(malloc(sizeof(_Ty) * n));
        if (size && container) { if (on_device()) aq::check(aqgpu_col_download(p, container, sizeof(_Ty) * size)); else memcpy((void*)p, (const void*)container, sizeof(_Ty) * size); }
        release_storage();
        container = p; capacity = n;
    }
    void release_storage() {
        if (capacity && container) { if (on_device()) aqgpu_col_free(container); else free(container); }
    }
    void release() { release_storage(); container = nullptr; size = capacity = 0; }
};

// ---- column statistics kept out of line (server/table.h:60-91 TableStats / populate_stats) -------------
// The reference embeds TableStats<T> in ColRef, which breaks get_col<i> for non-4-byte types
// (SURVEY.md section 7 "ABI"); here ColRef is 32 bytes for every T and statistics live in a side table.
namespace aq {
struct ColStats { long long lo, hi; };
inline std::unordered_map<const void*, ColStats>& col_stats() { static std::unordered_map<const void*, ColStats> m; return m; }
}

struct ColRef_cstorage { void* container; unsigned int size, capacity; const char* name; int ty; };   // server/table.h:34-39

template <class _Ty> class ColRef : public vector_type<_Ty> {
public:
    typedef ColRef<_Ty> Decayed_t;
    const char* name = "";
    types::Type_t ty = types::Types<_Ty>::getType();
    ColRef() : vector_type<_Ty>() {}
    ColRef(const uint32_t& n, const char* name = "") : vector_type<_Ty>(n), name(name) {}
    ColRef(const char* name) : vector_type<_Ty>(), name(name) {}
    ColRef(const uint32_t n, void* data, const char* name = "") : vector_type<_Ty>(n, data), name(name) {}     // table.h:101
    ColRef(const ColRef<_Ty>& o) : vector_type<_Ty>(static_cast<const vector_type<_Ty>&>(o)), name(o.name), ty(o.ty) {}
    ColRef(ColRef<_Ty>& o) : vector_type<_Ty>(static_cast<vector_type<_Ty>&>(o)), name(o.name), ty(o.ty) {}
    ColRef(ColRef<_Ty>&& o) noexcept : vector_type<_Ty>(std::move(static_cast<vector_type<_Ty>&>(o))), name(o.name), ty(o.ty) {}
    ColRef(vector_type<_Ty>&& v, const char* name = "") : vector_type<_Ty>(std::move(v)), name(name) {}
    ColRef<_Ty>& operator=(const ColRef<_Ty>& o) { vector_type<_Ty>::operator=(static_cast<const vector_type<_Ty>&>(o)); return *this; }
    ColRef<_Ty>& operator=(ColRef<_Ty>&& o) noexcept { vector_type<_Ty>::operator=(std::move(static_cast<vector_type<_Ty>&>(o))); return *this; }
    ColRef<_Ty>& operator=(vector_type<_Ty>&& o) noexcept { vector_type<_Ty>::operator=(std::move(o)); return *this; }
    void init(const char* nm = "") { this->container = nullptr; this->size = this->capacity = 0; name = nm; ty = types::Types<_Ty>::getType(); }
    void initfrom(uint32_t sz, void* data, const char* nm = "") { this->init_from(sz, data); name = nm; ty = types::Types<_Ty>::getType(); }
    template <class V> void initfrom(V&& v, const char* nm = "") {                                            // table.h:104-133
        this->init_from(v.size, v.container);
        if (!std::is_lvalue_reference<V>::value) { this->capacity = v.capacity; v.capacity = 0; }
        name = nm; ty = types::Types<_Ty>::getType();
    }
    ColRef<_Ty>* rename(const char* nm) { name = nm; return this; }
    ColRef_cstorage s() { return ColRef_cstorage{this->container, this->size, this->capacity, name, (int)ty}; }
    using vector_type<_Ty>::operator[];
    // statistics (ColRef::populate_stats, table.h:76-91): one streaming pass on the GPU
    bool populate_stats() {
        if constexpr (std::is_integral<_Ty>::value && (sizeof(_Ty) == 4 || sizeof(_Ty) == 8)) {
            if (aq::col_stats().count(this->container)) return true;
            long long lo = 0, hi = -1;
            aq::check(aqgpu_col_minmax(aq::dtype_of<_Ty>(), this->container, this->size, (int64_t*)&lo, (int64_t*)&hi));
            aq::col_stats()[this->container] = aq::ColStats{lo, hi};
            return true;
        }
        return false;
    }
    void set_stats(long long lo, long long hi) { aq::col_stats()[this->container] = aq::ColStats{lo, hi}; }
    // a copy of this column in the HBM column store
    ColRef<_Ty> to_device() const {
        ColRef<_Ty> d;
        if (this->on_device()) { d.init_from(this->size, this->container); d.name = name; return d; }
        void* p = nullptr;
        aq::check(aqgpu_col_alloc(sizeof(_Ty) * (this->size ? this->size : 1), &p));
        if (this->size) aq::check(aqgpu_col_upload(p, this->container, sizeof(_Ty) * this->size));
        d.container = static_cast<_Ty*>(p); d.size = this->size; d.capacity = this->size ? this->size : 1; d.name = name;
        return d;
    }
    // gather by an explicit index vector (table.h:184-189): out[j] = col[idx[j]], on the device; the result lives where the column lives
    ColRef<_Ty> operator[](const vector_type<uint32_t>& idx) const {
        ColRef<_Ty> out;
        if (this->on_device()) {
            void* p = nullptr;
            aq::check(aqgpu_col_alloc(sizeof(_Ty) * (idx.size ? idx.size : 1), &p));
            out.container = static_cast<_Ty*>(p); out.size = idx.size; out.capacity = idx.size ? idx.size : 1;
        } else out = ColRef<_Ty>(idx.size);
        if (idx.size) aq::check(aqgpu_take((int)sizeof(_Ty), this->container, this->size, idx.container, idx.size, out.container));
        return out;
    }
    // gather col[vecs[g]] (table.h:184-189): a proxy, evaluated by the aggregate that consumes it
    aq::GroupVec<_Ty> operator[](const aq::GroupRows& rows) const;
    template <class G, class = decltype(std::declval<const G&>().handle)> aq::GroupedCol<_Ty> operator[](const G& groups) const;
};
template <> class ColRef<void> : public ColRef<int> {};

template <class T> struct is_vector_type_impl : std::false_type {};
template <class T> struct is_vector_type_impl<vector_type<T>> : std::true_type {};
template <class T> struct is_vector_type_impl<ColRef<T>> : std::true_type {};

// ---- TableInfo (server/table.h:380-617) -----------------------------------------------------------------------
template <class... Types> struct TableInfo {
    const char* name;
    ColRef<void>* colrefs;
    uint32_t n_cols;
    typedef std::tuple<Types...> tuple_type;
    TableInfo(const char* name, const char** col_names) : name(name), n_cols(sizeof...(Types)) {
        colrefs = static_cast<ColRef<void>*>(malloc(sizeof(ColRef<void>) * (n_cols ? n_cols : 1)));
        for (uint32_t i = 0; i < n_cols; ++i) new (colrefs + i) ColRef<void>();
        init_cols(col_names, std::index_sequence_for<Types...>());
    }
    template <int i> auto& get_col() {
        static_assert(sizeof(ColRef<std::tuple_element_t<i, tuple_type>>) == sizeof(ColRef<void>), "one ColRef layout for every cell type");
        return *reinterpret_cast<ColRef<std::tuple_element_t<i, tuple_type>>*>(colrefs + i);
    }
    void print(const char* sep = ",", const char* end = "\n", uint32_t limit = 0xffffffffu) {
        print_impl(sep, end, limit, std::index_sequence_for<Types...>());
    }
    void printall(const char* sep = ",", const char* end = "\n", const char* = nullptr, FILE* = nullptr, uint32_t limit = 0xffffffffu) { print(sep, end, limit); }
    // in-place gather of every column by row ids (table.h:429-440); materialize_copy leaves this table alone (:441-445)
    void materialize(const vector_type<uint32_t>& idxs, TableInfo<Types...>* tbl = nullptr) {
        if (!tbl) tbl = this;
        materialize_impl(idxs, tbl, std::index_sequence_for<Types...>());
    }
    TableInfo<Types...>* materialize_copy(const vector_type<uint32_t>& idxs) {
        auto tbl = new TableInfo<Types...>(name, nullptr);
        for (uint32_t i = 0; i < n_cols; ++i) tbl->colrefs[i].name = colrefs[i].name;
        materialize(idxs, tbl);
        return tbl;
    }
    // row ids sorted by columns (table.h:446-461): cols >= 0 ascending by that column, cols < 0 descending by column -1 - cols;
    // `ord` (e.g. one group's row list, common/groupby.py:58) is sorted in place.  The reference uses std::sort, which leaves the
    // order of ties unspecified; here ties keep their current order (a stable sort: one of the orders the reference may produce).
    template <int... cols> vector_type<uint32_t>* order_by(vector_type<uint32_t>* ord = nullptr);
    // the distinct records (table.h:601-614).  The reference iterates an std::unordered_set (unspecified order); here the records
    // come in first-appearance order.
    TableInfo<Types...>* distinct();
private:
    template <size_t... I> void distinct_impl(std::index_sequence<I...>);
    template <size_t... I> void materialize_impl(const vector_type<uint32_t>& idxs, TableInfo<Types...>* tbl, std::index_sequence<I...>) {
        (void)std::initializer_list<int>{ (tbl->template get_col<(int)I>() = get_col<(int)I>()[idxs], 0)... };
    }
    template <size_t... I> void init_cols(const char** names, std::index_sequence<I...>) { ((get_col<(int)I>().init(names ? names[I] : "")), ...); }
    template <class C> static void print_one(const std::vector<C>& h, uint32_t r) {
        if constexpr (is_vector_type_impl<C>::value) h[r].out(8);
        else vector_type<C>::print_cell(h[r]);
    }
    template <size_t... I> void print_impl(const char* sep, const char* end, uint32_t limit, std::index_sequence<I...>) {
        auto hosts = std::make_tuple(host_cells(get_col<(int)I>())...);
        uint32_t rows = n_cols ? colrefs[0].size : 0;
        if (rows > limit) rows = limit;
        for (uint32_t i = 0; i < n_cols; ++i) std::cout << colrefs[i].name << (i + 1 < n_cols ? sep : end);
        for (uint32_t r = 0; r < rows; ++r) { ((print_one(std::get<I>(hosts), r), std::cout << (I + 1 < sizeof...(I) ? sep : end)), ...); }
    }
    template <class C> static std::vector<C> host_cells(ColRef<C>& c) {
        if constexpr (is_vector_type_impl<C>::value) { std::vector<C> h; h.reserve(c.size); for (uint32_t i = 0; i < c.size; ++i) h.emplace_back(c.container[i].container, c.container[i].size); return h; }
        else return c.to_host();
    }
};
template <class... Types> void print(TableInfo<Types...>& t, const char* sep = " ", const char* end = "\n") { t.print(sep, end); }

// ---- key records (server/types.h:356-450, server/hasher.h:66-140) -------------------------------------------
template <class... T> using record = std::tuple<T...>;
template <class T> struct decayS { typedef typename std::decay<T>::type type; };
template <class T> struct decayS<ColRef<T>> { typedef ColRef<T> type; };
template <class T> using decays = typename decayS<typename std::decay<T>::type>::type;
template <class T> struct value_type_impl { typedef T type; };
template <template <class...> class VT, class V0, class... V> struct value_type_impl<VT<V0, V...>> { typedef V0 type; };
template <class T> using value_type = typename value_type_impl<T>::type;
template <class... T> struct hasher {};                       // hashing happens on the GPU; kept for the template spelling
template <class Rec, template <class...> class H> struct transTypes_s { typedef H<Rec> type; };
template <class Rec, template <class...> class H> using transTypes = typename transTypes_s<Rec, H>::type;

// ---- group handle shared by the proxies -------------------------------------------------------------------------
namespace aq {
struct GroupRows {                 // `vecs[i]` of the generated loop: the row list of group i, unevaluated
    GroupsHandle* g; uint32_t gid; uint32_t size;     // .size: count(*) is emitted as val.size (engine/ast.py:763)
};
// A key column as the library sees it: one or two integer columns that are equal exactly when the reference's operator==
// holds (include/aqgpu.h "key columns that are not plain integers"; server/hasher.h:97-132).
struct LoweredKey {
    int kind = 0;                           // 0 plain / date image, AQ_KEY_TIME, AQ_KEY_TIMESTAMP, 4 = string
    int nparts = 1, first = 0;              // columns of the library's key tuple: [first, first + nparts)
    const void* p[2] = {nullptr, nullptr};
    int t[2] = {0, 0};
    std::vector<std::shared_ptr<void>> keep;        // device temporaries
    std::vector<uint32_t> dict_first;       // string: code -> first row
    const void* src = nullptr;              // string: the pointer array the codes index
};
inline std::shared_ptr<void> device_tmp(size_t bytes) {
    void* q = nullptr;
    check(aqgpu_col_alloc(bytes ? bytes : 16, &q));
    return std::shared_ptr<void>(q, [](void* x) { aqgpu_col_free(x); });
}
template <class T> inline LoweredKey lower_key(const T* col, uint32_t n) {
    LoweredKey k;
    if constexpr (std::is_same<T, types::date_t>::value) {
        k.p[0] = col; k.t[0] = AQ_UINT32;                                   // the 4-byte image is the key
    } else if constexpr (std::is_same<T, types::time_t>::value) {
        k.kind = AQ_KEY_TIME; k.t[0] = AQ_INT64;
        k.keep.push_back(device_tmp(size_t(n) * 8)); k.p[0] = k.keep[0].get();
        check(aqgpu_key_split(AQ_KEY_TIME, col, n, k.keep[0].get(), nullptr));
    } else if constexpr (std::is_same<T, types::timestamp_t>::value) {
        k.kind = AQ_KEY_TIMESTAMP; k.nparts = 2; k.t[0] = AQ_UINT32; k.t[1] = AQ_INT64;
        k.keep.push_back(device_tmp(size_t(n) * 4)); k.keep.push_back(device_tmp(size_t(n) * 8));
        k.p[0] = k.keep[0].get(); k.p[1] = k.keep[1].get();
        check(aqgpu_key_split(AQ_KEY_TIMESTAMP, col, n, k.keep[0].get(), k.keep[1].get()));
    } else if constexpr (std::is_same<T, astring_view>::value) {
        // ingest: the heap pointers mean nothing on the device -> pack the characters (Arrow layout), encode on the GPU
        k.kind = 4; k.t[0] = AQ_UINT32; k.src = col;
        std::vector<uint64_t> off(size_t(n) + 1, 0);
        for (uint32_t i = 0; i < n; ++i) off[i + 1] = off[i] + std::strlen(col[i].rstr);
        std::vector<char> bytes(off[n] ? off[n] : 1);
        for (uint32_t i = 0; i < n; ++i) std::memcpy(bytes.data() + off[i], col[i].rstr, off[i + 1] - off[i]);
        k.keep.push_back(device_tmp(size_t(n) * 4)); k.p[0] = k.keep[0].get();
        k.dict_first.resize(n ? n : 1);
        uint32_t nd = 0;
        check(aqgpu_string_dict(bytes.data(), off[n], off.data(), n, static_cast<uint32_t*>(k.keep[0].get()), n, k.dict_first.data(), &nd));
        k.dict_first.resize(nd);
    } else {
        static_assert(dtype_of<T>() >= 0, "key columns: int32 / uint32 / int64, date / time / timestamp, astring_view");
        k.p[0] = col; k.t[0] = dtype_of<T>();
    }
    return k;
}

struct GroupsHandle {
    aqgpu_groups_t* h = nullptr;
    uint32_t nrows = 0, ngroups = 0;
    std::vector<LoweredKey> lowered;                  // one per key column of the query
    // key column i of the result, ngroups cells of K, first-appearance order
    template <class K> void fetch_key(int i, K* out) {
        const LoweredKey& k = lowered[i];
        if (!ngroups) return;
        if (k.kind == AQ_KEY_TIMESTAMP) {
            std::vector<uint32_t> a(ngroups); std::vector<long long> b(ngroups);
            check(aqgpu_groups_keys(h, k.first, a.data()));
            check(aqgpu_groups_keys(h, k.first + 1, b.data()));
            check(aqgpu_key_join(AQ_KEY_TIMESTAMP, a.data(), b.data(), ngroups, out));
        } else if (k.kind == 4) {
            if constexpr (std::is_same<K, astring_view>::value) {
                std::vector<uint32_t> code(ngroups);
                check(aqgpu_groups_keys(h, k.first, code.data()));
                const astring_view* src = static_cast<const astring_view*>(k.src);
                for (uint32_t g = 0; g < ngroups; ++g) out[g] = src[k.dict_first[code[g]]];
            }
        } else {
            check(aqgpu_groups_keys(h, k.first, out));            // plain cells; date / time images are valid cells
        }
    }
    std::vector<uint32_t> offsets;                    // host copy of ht_base + terminal (ngroups + 1)
    std::vector<GroupRows> rows;                      // backing store of HashTableComponents::values
    std::shared_ptr<void> keys;                       // backing store of HashTableComponents::keys
    std::map<std::tuple<int, const void*, const void*, uint32_t>, std::shared_ptr<void>> memo;
    ~GroupsHandle() { if (h) aqgpu_groups_free(h); }
    void fetch_offsets() {
        if (!offsets.empty()) return;
        offsets.assign(size_t(ngroups) + 1, 0u);
        check(aqgpu_groups_csr(h, offsets.data(), nullptr));
        rows.resize(ngroups);
        for (uint32_t i = 0; i < ngroups; ++i) rows[i] = GroupRows{this, i, offsets[i + 1] - offsets[i]};
    }
    // op(v) for ALL groups, evaluated once on the GPU, kept as a host column of ngroups cells of R
    template <class R> const R* reduce_all(int op, int vtype, const void* v, const void* v2 = nullptr) {
        auto key = std::make_tuple(op, v, v2, 0u);
        auto it = memo.find(key);
        if (it == memo.end()) {
            std::shared_ptr<void> buf(malloc(sizeof(R) * (ngroups ? ngroups : 1)), free);
            check(aqgpu_group_reduce(h, op, vtype, v, v2, dtype_of<R>(), buf.get()));
            it = memo.emplace(key, buf).first;
        }
        return static_cast<const R*>(it->second.get());
    }
    // scan op(w, v) for ALL groups: nrows cells of R in the CSR order of the row lists
    template <class R> const R* scan_all(int op, uint32_t w, int vtype, const void* v) {
        auto key = std::make_tuple(100 + op, v, (const void*)nullptr, w);
        auto it = memo.find(key);
        if (it == memo.end()) {
            std::shared_ptr<void> buf(malloc(sizeof(R) * (nrows ? nrows : 1)), free);
            check(aqgpu_group_scan(h, op, w, vtype, v, dtype_of<R>(), buf.get()));
            it = memo.emplace(key, buf).first;
        }
        return static_cast<const R*>(it->second.get());
    }
    template <class T> const T* gather_all(const T* v) {          // col[vecs[g]] for all groups, CSR order
        auto key = std::make_tuple(200, (const void*)v, (const void*)nullptr, 0u);
        auto it = memo.find(key);
        if (it == memo.end()) {
            std::shared_ptr<void> buf(malloc(sizeof(T) * (nrows ? nrows : 1)), free);
            check(aqgpu_group_gather(h, dtype_of<T>(), v, buf.get()));
            it = memo.emplace(key, buf).first;
        }
        return static_cast<const T*>(it->second.get());
    }
};
template <class T> struct GroupVec {       // col[vecs[i]]: the values of one group, unevaluated
    const T* col; GroupsHandle* g; uint32_t gid; uint32_t size;
    operator vector_type<T>() const {      // materialised only when generated code does arithmetic on it
        g->fetch_offsets();
        return vector_type<T>(const_cast<T*>(g->gather_all(col)) + g->offsets[gid], size);
    }
    // Host-visible cells for user code that walks a group itself (elemental-return UDFs, engine/ast.py:768-777,
    // tests/udf3.a `x[i]`, `x(0, i)`): the gather col[vecs[g]] runs ONCE on the device for all groups (aqgpu_group_gather), the
    // UDF body then reads its group's slice of the host copy.
    mutable const T* host = nullptr;       // this group's slice of the gathered host copy, resolved on first use
    const T* data() const { if (!host) { g->fetch_offsets(); host = g->gather_all(col) + g->offsets[gid]; } return host; }
    const T& operator[](uint32_t i) const { return data()[i]; }
    const T* begin() const { return data(); }
    const T* end() const { return data() + size; }
    vector_type<T> subvec(uint32_t start, uint32_t stop) const { return vector_type<T>(const_cast<T*>(data()) + start, stop - start); }   // vector_type.hpp:353-355
    vector_type<T> operator()(uint32_t start, uint32_t stop) const { return subvec(start, stop); }
};
template <class T> struct GroupedCol {     // col[g]: a column seen through ALL groups (whole-column code)
    const T* col; uint32_t n; std::shared_ptr<GroupsHandle> handle;
};

template <class... Ts, size_t... I>
inline std::shared_ptr<GroupsHandle> build_groups(std::index_sequence<I...>, int row_order, const ColRef<Ts>&... cols) {
    constexpr int nk = sizeof...(Ts);
    static_assert(nk >= 1, "group by needs at least one key column");
    const uint32_t n = std::get<0>(std::tie(cols...)).size;
    auto g = std::make_shared<GroupsHandle>();
    g->lowered = { lower_key<Ts>(cols.container, n)... };
    const void* kp[2 * nk]; int kt[2 * nk];
    int64_t lo[2 * nk], hi[2 * nk];
    bool have = true;
    double prod = 1.0;
    int m = 0;
    for (auto& k : g->lowered) {
        k.first = m;
        for (int j = 0; j < k.nparts; ++j, ++m) {
            kp[m] = k.p[j]; kt[m] = k.t[j];
            auto it = k.kind == 0 ? col_stats().find(k.p[j]) : col_stats().end();
            if (it == col_stats().end()) have = false; else { lo[m] = it->second.lo; hi[m] = it->second.hi; prod *= double(hi[m] - lo[m] + 1); }
        }
    }
    const uint32_t hint = have ? (uint32_t)std::min<double>(prod, double(n)) : 0u;
    check(aqgpu_groupby_build(kp, kt, m, n, have ? lo : nullptr, have ? hi : nullptr, row_order, hint, &g->h));
    g->nrows = n;
    check(aqgpu_groups_size(g->h, &g->ngroups));
    return g;
}

// Whole-column group-by: keys stay in HBM until asked for.
template <class... Ts> struct Groups {
    std::shared_ptr<GroupsHandle> handle;
    uint32_t size = 0;
    template <int i> ColRef<std::tuple_element_t<i, std::tuple<Ts...>>> key() const {
        typedef std::tuple_element_t<i, std::tuple<Ts...>> K;
        ColRef<K> out(handle->ngroups);
        handle->template fetch_key<K>(i, out.container);
        return out;
    }
    ColRef<uint32_t> offsets() const { handle->fetch_offsets(); ColRef<uint32_t> o((uint32_t)handle->offsets.size()); std::copy(handle->offsets.begin(), handle->offsets.end(), o.container); return o; }
    ColRef<uint32_t> counts() const {
        ColRef<uint32_t> c(handle->ngroups);
        if (handle->ngroups) check(aqgpu_group_reduce(handle->h, AQ_COUNT, AQ_INT32, nullptr, nullptr, AQ_UINT32, c.container));
        return c;
    }
};
template <class... Ts> Groups<Ts...> group_by(const ColRef<Ts>&... cols) {
    Groups<Ts...> g;
    g.handle = build_groups<Ts...>(std::index_sequence_for<Ts...>(), AQ_ROWS_DESC, cols...);
    g.size = g.handle->ngroups;
    return g;
}
}  // namespace aq

template <class _Ty> aq::GroupVec<_Ty> ColRef<_Ty>::operator[](const aq::GroupRows& r) const { return aq::GroupVec<_Ty>{this->container, r.g, r.gid, r.size}; }
template <class _Ty> template <class G, class> aq::GroupedCol<_Ty> ColRef<_Ty>::operator[](const G& g) const { return aq::GroupedCol<_Ty>{this->container, this->size, g.handle}; }

// ---- aggregates of a single value (server/aggregations.h:498-527): `SELECT max(x)` where x is a scalar expression ------------
#define AQ_SCALAR(NAME, RET) template <class T, class = std::enable_if_t<std::is_arithmetic<T>::value>> constexpr T NAME(const T& v) { (void)v; return RET; }
#define AQ_SCALARW(NAME, RET) template <class T, class = std::enable_if_t<std::is_arithmetic<T>::value>> constexpr T NAME(uint32_t, const T& v) { (void)v; return RET; }
AQ_SCALAR(count, 1) AQ_SCALAR(var, 0) AQ_SCALAR(vars, 0) AQ_SCALAR(stddev, 0) AQ_SCALAR(stddevs, 0)
AQ_SCALAR(max, v) AQ_SCALAR(min, v) AQ_SCALAR(avg, v) AQ_SCALAR(sum, v) AQ_SCALAR(maxs, v) AQ_SCALAR(mins, v) AQ_SCALAR(avgs, v) AQ_SCALAR(sums, v)
AQ_SCALAR(last, v) AQ_SCALAR(first, v) AQ_SCALAR(prev, v) AQ_SCALAR(aggnext, v) AQ_SCALAR(deltas, 0) AQ_SCALAR(ratios, 1)
AQ_SCALARW(varw, 0) AQ_SCALARW(stddevw, 0) AQ_SCALARW(maxw, v) AQ_SCALARW(minw, v) AQ_SCALARW(avgw, v) AQ_SCALARW(sumw, v) AQ_SCALARW(ratiow, 1)
#undef AQ_SCALAR
#undef AQ_SCALARW

// ---- host-visible views for user modules ---------------------------------------------------------------
// SDK modules take ColRef<T> / vector_type<T> BY VALUE and dereference `container` on the CPU (sdk/example.cpp:4,
// sdk/aquery.h:105-128 ColRef_storage); a column of the HBM store is materialised to host memory at that call site only.
namespace aq {
template <class T> ColRef<T> host_view(const ColRef<T>& c) {
    ColRef<T> h;
    if (!c.on_device()) { h.init_from(c.size, c.container); h.name = c.name; return h; }       // already dereferenceable: a view
    h = ColRef<T>(c.size, c.name);
    if (c.size) check(aqgpu_col_download(h.container, c.container, sizeof(T) * c.size));
    return h;
}
}  // namespace aq

// ---- TableInfo::order_by / distinct (server/table.h:446-461, 601-614) on the device --------------------------------------
template <class... Types> template <int... cols>
vector_type<uint32_t>* TableInfo<Types...>::order_by(vector_type<uint32_t>* ord) {
    constexpr int nc = sizeof...(cols);
    static_assert(nc >= 1, "order_by needs at least one column");
    const uint32_t rows = n_cols ? colrefs[0].size : 0;
    const bool subset = ord != nullptr;
    const uint32_t m = subset ? ord->size : rows;
    if (!ord) ord = new vector_type<uint32_t>(rows);
    if (m == 0) return ord;
    // the sort columns, restricted to the rows of `ord` when a row list is given
    const void* cp[nc]; int ct[nc]; const int desc[nc] = { (cols < 0 ? 1 : 0)... };
    std::vector<std::shared_ptr<void>> keep;
    int k = 0;
    auto bind = [&](auto& col) {
        typedef typename std::decay_t<decltype(col)>::value_t C;
        static_assert(aq::dtype_of<C>() >= 0, "order_by: integer / float columns");
        ct[k] = aq::dtype_of<C>();
        if (subset) {
            auto tmp = std::make_shared<ColRef<C>>(col[*ord]);
            cp[k] = tmp->container; keep.push_back(tmp);
        } else cp[k] = col.container;
        ++k;
    };
    (void)std::initializer_list<int>{ (bind(get_col<(cols >= 0 ? cols : -1 - cols)>()), 0)... };
    // one group holding every row in table order, its row list re-ordered by the segmented radix sort
    std::vector<uint32_t> zero(m, 0u);
    const void* kp[1] = { zero.data() }; const int kt[1] = { AQ_UINT32 }; const int64_t lo[1] = {0}, hi[1] = {0};
    aqgpu_groups_t* h = nullptr;
    aq::check(aqgpu_groupby_build(kp, kt, 1, m, lo, hi, AQ_ROWS_ASC, 1, &h));
    std::shared_ptr<void> guard(h, [](void* x) { aqgpu_groups_free(static_cast<aqgpu_groups_t*>(x)); });
    aq::check(aqgpu_groups_order_by(h, cp, ct, desc, nc));
    std::vector<uint32_t> perm(m), off(2);
    aq::check(aqgpu_groups_csr(h, off.data(), perm.data()));
    if (subset) { std::vector<uint32_t> old(ord->container, ord->container + m); for (uint32_t j = 0; j < m; ++j) ord->container[j] = old[perm[j]]; }
    else std::copy(perm.begin(), perm.end(), ord->container);
    return ord;
}
template <class... Types> template <size_t... I> void TableInfo<Types...>::distinct_impl(std::index_sequence<I...>) {
    auto g = aq::group_by(get_col<(int)I>()...);
    auto keys = std::make_tuple(g.template key<(int)I>()...);              // fetched before any column is replaced
    (void)std::initializer_list<int>{ (get_col<(int)I>() = std::move(std::get<I>(keys)), 0)... };
}
template <class... Types> TableInfo<Types...>* TableInfo<Types...>::distinct() {
    if (n_cols && colrefs[0].size) distinct_impl(std::index_sequence_for<Types...>());
    return this;
}

// ---- HashTableComponents / HashTableFactory / AQHashTable (server/hasher.h:146-357) --------------------------------
template <class... Ty> struct HashTableComponents {
    uint32_t size;
    std::vector<std::tuple<Ty...>>* keys;
    aq::GroupRows* values;                 // reference: vector_type<uint32_t>*; here proxies of the row lists
    uint32_t* offsets;
    std::shared_ptr<aq::GroupsHandle> handle;     // extension: keeps the HBM-resident groups alive
};

namespace aq {
template <class... Ts, size_t... I>
inline HashTableComponents<Ts...> make_components(std::index_sequence<I...>, std::shared_ptr<GroupsHandle> g) {
    g->fetch_offsets();
    auto keys = std::make_shared<std::vector<std::tuple<Ts...>>>(g->ngroups);
    std::tuple<std::vector<Ts>...> cols{ std::vector<Ts>(g->ngroups)... };
    (g->template fetch_key<Ts>((int)I, std::get<I>(cols).data()), ...);
    for (uint32_t i = 0; i < g->ngroups; ++i) (*keys)[i] = std::make_tuple(std::get<I>(cols)[i]...);
    g->keys = keys;
    return HashTableComponents<Ts...>{g->ngroups, keys.get(), g->rows.data(), g->offsets.data(), g};
}
}  // namespace aq

template <class Key, class Hash = void> struct HashTableFactory {
    // the emitter spells explicit template arguments (engine/ast.py:668-670); they are accepted and ignored
    template <class... Ignored, class... Ts> static HashTableComponents<Ts...> get(const ColRef<Ts>&... cols) {
        auto g = aq::build_groups<Ts...>(std::index_sequence_for<Ts...>(), AQ_ROWS_DESC, cols...);
        return aq::make_components<Ts...>(std::index_sequence_for<Ts...>(), g);
    }
};

// The previous generation of the emitted code (engine/ast.py:664-676, mem_opt.cpp:31-52):
//   AQHashTable<rec_t, ...> g{total}; g.hashtable_push_all<...>(cols..., total) | g.hashtable_push(key, i) per row;
//   vecs = g.ht_postproc(total); g.size(); g.values()[i]; g.ht_base
template <class Key, class Hash = void> class AQHashTable;
template <class... Ks, class Hash> class AQHashTable<std::tuple<Ks...>, Hash> {
public:
    uint32_t *reversemap = nullptr, *mapbase = nullptr, *ht_base = nullptr;
    AQHashTable() = default;
    explicit AQHashTable(uint32_t sz) { init(sz); }
    void init(uint32_t sz) { std::apply([&](auto&... v) { (v.reserve(sz), ...); }, pushed); }
    template <class... Ignored> void hashtable_push_all(const ColRef<Ks>&... cols, uint32_t len) {
        (void)len;
        g = aq::build_groups<Ks...>(std::index_sequence_for<Ks...>(), AQ_ROWS_DESC, cols...);
    }
    template <class Tup> void hashtable_push(Tup&& k, uint32_t) { push_impl(k, std::index_sequence_for<Ks...>()); }   // key tuple of row i
    aq::GroupRows* ht_postproc(uint32_t sz) {
        if (!g) {          // per-row pushes: the buffered key columns are grouped now, in one call
            g = std::apply([&](auto&... v) {
                return aq::build_groups<Ks...>(std::index_sequence_for<Ks...>(), AQ_ROWS_DESC, ColRef<Ks>((uint32_t)v.size(), v.data())...); }, pushed);
        }
        (void)sz;
        comp = aq::make_components<Ks...>(std::index_sequence_for<Ks...>(), g);
        ht_base = comp.offsets;
        return comp.values;
    }
    uint32_t size() const { return g ? g->ngroups : 0; }
    std::vector<std::tuple<Ks...>>& values() { return *comp.keys; }
    std::shared_ptr<aq::GroupsHandle> handle() const { return g; }
private:
    template <class Tup, size_t... I> void push_impl(const Tup& k, std::index_sequence<I...>) { (std::get<I>(pushed).push_back(std::get<I>(k)), ...); }
    std::tuple<std::vector<Ks>...> pushed;
    std::shared_ptr<aq::GroupsHandle> g;
    HashTableComponents<Ks...> comp{};
};

// ---- aggregates (server/aggregations.h) -------------------------------------------------------------------------
// (1) per-group proxies: the value for group gid of a column evaluated once for all groups
template <class T> types::GetLongType<T> sum(const aq::GroupVec<T>& v) { return v.g->template reduce_all<types::GetLongType<T>>(AQ_SUM, aq::dtype_of<T>(), v.col)[v.gid]; }   // :19-27
template <class T> double avg(const aq::GroupVec<T>& v) { return v.g->template reduce_all<double>(AQ_AVG, aq::dtype_of<T>(), v.col)[v.gid]; }                              // :28-32
template <class T> T max(const aq::GroupVec<T>& v) { return v.g->template reduce_all<T>(AQ_MAX, aq::dtype_of<T>(), v.col)[v.gid]; }                                       // :71-78
template <class T> T min(const aq::GroupVec<T>& v) { return v.g->template reduce_all<T>(AQ_MIN, aq::dtype_of<T>(), v.col)[v.gid]; }                                       // :79-86
template <class T> T first(const aq::GroupVec<T>& v) { return v.g->template reduce_all<T>(AQ_FIRST, aq::dtype_of<T>(), v.col)[v.gid]; }                                   // :493-497
template <class T> T last(const aq::GroupVec<T>& v) { return v.g->template reduce_all<T>(AQ_LAST, aq::dtype_of<T>(), v.col)[v.gid]; }                                     // :487-491
template <class T> uint32_t count(const aq::GroupVec<T>& v) { return v.size; }                                                                                          // :10-16
inline uint32_t count(const aq::GroupRows& r) { return r.size; }
template <class T> double var(const aq::GroupVec<T>& v) { return v.g->template reduce_all<double>(AQ_VAR, aq::dtype_of<T>(), v.col)[v.gid]; }                              // :332-348
template <class T> double stddev(const aq::GroupVec<T>& v) { return v.g->template reduce_all<double>(AQ_STDDEV, aq::dtype_of<T>(), v.col)[v.gid]; }                        // :413-415
template <class T> double corr(const aq::GroupVec<T>& a, const aq::GroupVec<T>& b) { return a.g->template reduce_all<double>(AQ_CORR, aq::dtype_of<T>(), a.col, b.col)[a.gid]; }   // :383-407

namespace aq {
template <class R, class T> vector_type<R> scan_view(const GroupVec<T>& v, int op, uint32_t w) {
    v.g->fetch_offsets();
    const R* flat = v.g->template scan_all<R>(op, w, dtype_of<T>(), v.col);
    return vector_type<R>(const_cast<R*>(flat) + v.g->offsets[v.gid], v.size);
}
template <class R, class T, class Ret> void scan_into(const GroupVec<T>& v, int op, uint32_t w, Ret& ret) {
    const vector_type<R> s = scan_view<R, T>(v, op, w);
    if (ret.size < s.size) ret.resize(s.size);
    for (uint32_t i = 0; i < s.size; ++i) ret.container[i] = s.container[i];
}
}  // namespace aq
// running / windowed / neighbour aggregates: 1-argument form returns the group's slice of the memoised
// CSR result, the "write into ret" form (engine/expr.py:373-385, mem_opt.cpp:60) copies it into the cell
#define AQ_SCAN0(NAME, OP, RT)                                                                                      \
    template <class T> vector_type<RT> NAME(const aq::GroupVec<T>& v) { return aq::scan_view<RT, T>(v, OP, 0); }      \
    template <class T, class Ret> void NAME(const aq::GroupVec<T>& v, Ret& ret) { aq::scan_into<RT, T>(v, OP, 0, ret); }
#define AQ_SCANW(NAME, OP, RT)                                                                                      \
    template <class T> vector_type<RT> NAME(uint32_t w, const aq::GroupVec<T>& v) { return aq::scan_view<RT, T>(v, OP, w); } \
    template <class T, class Ret> void NAME(uint32_t w, const aq::GroupVec<T>& v, Ret& ret) { aq::scan_into<RT, T>(v, OP, w, ret); }
AQ_SCAN0(sums, AQ_SUMS, types::GetLongType<T>)      // :203-217
AQ_SCAN0(avgs, AQ_AVGS, double)                     // :219-236
AQ_SCAN0(mins, AQ_MINS, T)                          // :89-106
AQ_SCAN0(maxs, AQ_MAXS, T)                          // :108-125
AQ_SCAN0(deltas, AQ_DELTAS, T)                      // :439-453
AQ_SCAN0(prev, AQ_PREV, T)                          // :455-469
AQ_SCAN0(aggnext, AQ_NEXT, T)                       // :471-485
AQ_SCANW(sumw, AQ_SUMW, types::GetLongType<T>)      // :238-256
AQ_SCANW(avgw, AQ_AVGW, double)                     // :258-281
AQ_SCANW(minw, AQ_MINW, T)                          // :127-146
AQ_SCANW(maxw, AQ_MAXW, T)                          // :148-167
AQ_SCANW(ratiow, AQ_RATIOW, types::GetFPType<T>)    // :169-201
AQ_SCAN0(vars, AQ_VARS, double)                     // :350-381
AQ_SCAN0(stddevs, AQ_STDDEVS, double)               // :419-425
AQ_SCANW(varw, AQ_VARW, double)                     // :283-330 (see include/aqgpu.h: the reference's recurrence is undefined at i == w)
AQ_SCANW(stddevw, AQ_STDDEVW, double)               // :427-436
template <class T> vector_type<types::GetFPType<T>> ratios(const aq::GroupVec<T>& v) { return aq::scan_view<types::GetFPType<T>, T>(v, AQ_RATIOW, 1); }
// SQL spellings with a window argument (common/types.py:285-290, 349-352): sums(w, x) == sumw(w, x) ...
template <class T> auto sums(uint32_t w, const aq::GroupVec<T>& v) { return sumw(w, v); }
template <class T> auto vars(uint32_t w, const aq::GroupVec<T>& v) { return varw(w, v); }
template <class T> auto stddevs(uint32_t w, const aq::GroupVec<T>& v) { return stddevw(w, v); }
template <class T> auto avgs(uint32_t w, const aq::GroupVec<T>& v) { return avgw(w, v); }
template <class T> auto mins(uint32_t w, const aq::GroupVec<T>& v) { return minw(w, v); }
template <class T> auto maxs(uint32_t w, const aq::GroupVec<T>& v) { return maxw(w, v); }

// (2) whole-column forms: result columns for all groups in one call.  The result lives where the input
// column lives (HBM for device-resident inputs, host otherwise).
namespace aq {
template <class R> ColRef<R> alloc_like_input(const void* input, uint32_t n) {
    ColRef<R> out;
    if (input && aqgpu_ptr_is_device(input) == 1) {
        void* p = nullptr;
        check(aqgpu_col_alloc(sizeof(R) * (n ? n : 1), &p));
        out.container = static_cast<R*>(p); out.size = n; out.capacity = n ? n : 1;
    } else out = ColRef<R>(n);
    return out;
}
template <class R, class T> ColRef<R> reduce_col(const GroupedCol<T>& c, int op, const T* v2 = nullptr) {
    ColRef<R> out = alloc_like_input<R>(c.col, c.handle->ngroups);
    if (c.handle->ngroups) check(aqgpu_group_reduce(c.handle->h, op, dtype_of<T>(), c.col, v2, dtype_of<R>(), out.container));
    return out;
}
// vector-valued result: flat values in the CSR order of the row lists + the groups' offsets (engine/ast.py:696-718)
template <class R> struct VecCol {
    ColRef<R> flat;
    std::shared_ptr<GroupsHandle> handle;
    // per-group views for a TableInfo column of vector_type<R> cells (host-resident results only)
    ColRef<vector_type<R>> views() {
        handle->fetch_offsets();
        ColRef<vector_type<R>> v(handle->ngroups);
        for (uint32_t i = 0; i < handle->ngroups; ++i) new (v.container + i) vector_type<R>(flat.container + handle->offsets[i], handle->offsets[i + 1] - handle->offsets[i]);
        return v;
    }
};
template <class R, class T> VecCol<R> scan_col(const GroupedCol<T>& c, int op, uint32_t w) {
    VecCol<R> out{alloc_like_input<R>(c.col, c.handle->nrows), c.handle};
    if (c.handle->nrows) check(aqgpu_group_scan(c.handle->h, op, w, dtype_of<T>(), c.col, dtype_of<R>(), out.flat.container));
    return out;
}
}  // namespace aq
template <class T> ColRef<types::GetLongType<T>> sum(const aq::GroupedCol<T>& c) { return aq::reduce_col<types::GetLongType<T>, T>(c, AQ_SUM); }
template <class T> ColRef<double> avg(const aq::GroupedCol<T>& c) { return aq::reduce_col<double, T>(c, AQ_AVG); }
template <class T> ColRef<T> max(const aq::GroupedCol<T>& c) { return aq::reduce_col<T, T>(c, AQ_MAX); }
template <class T> ColRef<T> min(const aq::GroupedCol<T>& c) { return aq::reduce_col<T, T>(c, AQ_MIN); }
template <class T> ColRef<T> first(const aq::GroupedCol<T>& c) { return aq::reduce_col<T, T>(c, AQ_FIRST); }
template <class T> ColRef<T> last(const aq::GroupedCol<T>& c) { return aq::reduce_col<T, T>(c, AQ_LAST); }
template <class T> ColRef<uint32_t> count(const aq::GroupedCol<T>& c) { return aq::reduce_col<uint32_t, T>(c, AQ_COUNT); }
template <class T> ColRef<double> var(const aq::GroupedCol<T>& c) { return aq::reduce_col<double, T>(c, AQ_VAR); }
template <class T> ColRef<double> stddev(const aq::GroupedCol<T>& c) { return aq::reduce_col<double, T>(c, AQ_STDDEV); }
template <class T> ColRef<double> corr(const aq::GroupedCol<T>& a, const aq::GroupedCol<T>& b) { return aq::reduce_col<double, T>(a, AQ_CORR, b.col); }
#define AQ_WSCAN0(NAME, OP, RT) template <class T> aq::VecCol<RT> NAME(const aq::GroupedCol<T>& c) { return aq::scan_col<RT, T>(c, OP, 0); }
#define AQ_WSCANW(NAME, OP, RT) template <class T> aq::VecCol<RT> NAME(uint32_t w, const aq::GroupedCol<T>& c) { return aq::scan_col<RT, T>(c, OP, w); }
AQ_WSCAN0(sums, AQ_SUMS, types::GetLongType<T>) AQ_WSCAN0(avgs, AQ_AVGS, double) AQ_WSCAN0(mins, AQ_MINS, T) AQ_WSCAN0(maxs, AQ_MAXS, T)
AQ_WSCAN0(deltas, AQ_DELTAS, T) AQ_WSCAN0(prev, AQ_PREV, T) AQ_WSCAN0(aggnext, AQ_NEXT, T)
AQ_WSCANW(sumw, AQ_SUMW, types::GetLongType<T>) AQ_WSCANW(avgw, AQ_AVGW, double) AQ_WSCANW(minw, AQ_MINW, T) AQ_WSCANW(maxw, AQ_MAXW, T)
AQ_WSCANW(ratiow, AQ_RATIOW, types::GetFPType<T>)
AQ_WSCAN0(vars, AQ_VARS, double) AQ_WSCAN0(stddevs, AQ_STDDEVS, double) AQ_WSCANW(varw, AQ_VARW, double) AQ_WSCANW(stddevw, AQ_STDDEVW, double)
AQ_WSCANW(vars, AQ_VARW, double) AQ_WSCANW(stddevs, AQ_STDDEVW, double)      // SQL: stddevs(3, price) (benchmark/quries/Aquery/q10.a)
AQ_WSCANW(sums, AQ_SUMW, types::GetLongType<T>) AQ_WSCANW(avgs, AQ_AVGW, double) AQ_WSCANW(mins, AQ_MINW, T) AQ_WSCANW(maxs, AQ_MAXW, T)
template <class T> aq::VecCol<types::GetFPType<T>> ratios(const aq::GroupedCol<T>& c) { return aq::scan_col<types::GetFPType<T>, T>(c, AQ_RATIOW, 1); }

// (3) non-grouped forms: the whole column is ONE group in table order -- `SELECT avgs(3, sales) FROM sale ASSUMING ASC month`
// (tests/moving_avg.a:7; the vector functions of server/aggregations.h applied to a plain ColRef).  One row per input row.
namespace aq {
template <class T> GroupedCol<T> whole_column(const ColRef<T>& c) {
    auto g = std::make_shared<GroupsHandle>();
    std::vector<uint32_t> zero(c.size ? c.size : 1, 0u);
    const void* kp[1] = { zero.data() };
    const int kt[1] = { AQ_UINT32 };
    const int64_t lo[1] = {0}, hi[1] = {0};
    check(aqgpu_groupby_build(kp, kt, 1, c.size, lo, hi, AQ_ROWS_ASC, 1, &g->h));
    g->nrows = c.size;
    check(aqgpu_groups_size(g->h, &g->ngroups));
    return GroupedCol<T>{c.container, c.size, g};
}
}  // namespace aq
#define AQ_PSCAN0(NAME) template <class T> auto NAME(const ColRef<T>& c) { return NAME(aq::whole_column(c)).flat; }
#define AQ_PSCANW(NAME) template <class T> auto NAME(uint32_t w, const ColRef<T>& c) { return NAME(w, aq::whole_column(c)).flat; }
AQ_PSCAN0(sums) AQ_PSCAN0(avgs) AQ_PSCAN0(mins) AQ_PSCAN0(maxs) AQ_PSCAN0(deltas) AQ_PSCAN0(prev) AQ_PSCAN0(aggnext) AQ_PSCAN0(ratios)
AQ_PSCAN0(vars) AQ_PSCAN0(stddevs)
AQ_PSCANW(sums) AQ_PSCANW(avgs) AQ_PSCANW(mins) AQ_PSCANW(maxs) AQ_PSCANW(sumw) AQ_PSCANW(avgw) AQ_PSCANW(minw) AQ_PSCANW(maxw) AQ_PSCANW(ratiow)
AQ_PSCANW(vars) AQ_PSCANW(stddevs) AQ_PSCANW(varw) AQ_PSCANW(stddevw)
#undef AQ_PSCAN0
#undef AQ_PSCANW
// scalar aggregates of a whole column (aggregations.h:10-86, 332-415): one cell
#define AQ_PRED(NAME, RT) template <class T> RT NAME(const ColRef<T>& c) { auto r = NAME(aq::whole_column(c)); return c.size ? aq::cell0(r) : RT{}; }
namespace aq {
template <class R> R cell0(const ColRef<R>& r) {          // first cell of a result column that lives where the input lives
    R v{};
    if (r.size) { if (aqgpu_ptr_is_device(r.container) == 1) check(aqgpu_col_download(&v, r.container, sizeof(R))); else v = r.container[0]; }
    return v;
}
}  // namespace aq
AQ_PRED(sum, types::GetLongType<T>) AQ_PRED(avg, double) AQ_PRED(max, T) AQ_PRED(min, T) AQ_PRED(var, double) AQ_PRED(stddev, double)
#undef AQ_PRED
#undef AQ_SCAN0
#undef AQ_SCANW
#undef AQ_WSCAN0
#undef AQ_WSCANW

// ---- expressions over result columns (server/table.h:820-973 vector arithmetic, aggregations.h:34-69,409-411) -----
// Results of whole-column aggregates are columns of ngroups cells that live where their inputs live; the
// operators the generated code applies to them run on the device for device-resident operands.
template <class T> ColRef<T> operator-(const ColRef<T>& a, const ColRef<T>& b) {                    // q7: max(v1) - min(v2)
    ColRef<T> out = aq::alloc_like_input<T>(a.container, a.size);
    if (a.size) aq::check(aqgpu_vec_sub(aq::dtype_of<T>(), a.container, b.container, out.container, a.size));
    return out;
}
// General forms (server/vector_type.hpp:372-396, server/table.h:826-973): `a op b` for columns of different element types,
// column op scalar, scalar op column; comparisons give ColRef<bool>.  Result element types are the reference's
// (aqgpu_vec_out_type mirrors its templates; tests/test_gpu_vecops.py pins the table to the reference build).
namespace aq {
template <class A, class B> struct coercion {               // types::Coercion<T1, T2> for {int, long long, float, double} (types.h:264-275)
    static constexpr bool fp = std::is_floating_point<A>::value || std::is_floating_point<B>::value;
    static constexpr bool wide = sizeof(A) == 8 || sizeof(B) == 8;
    typedef std::conditional_t<fp, std::conditional_t<wide, double, float>, std::conditional_t<wide, long long, int>> type;
};
template <class T> struct is_vec_cell : std::integral_constant<bool, std::is_same<T, int>::value || std::is_same<T, long long>::value ||
    std::is_same<T, long>::value || std::is_same<T, float>::value || std::is_same<T, double>::value || std::is_same<T, bool>::value> {};
template <class T> using cell_t = std::conditional_t<std::is_same<T, long>::value, long long, std::conditional_t<std::is_same<T, bool>::value, int, T>>;
template <class R, class A, class B>
ColRef<R> vec_op(int op, const A* a, uint32_t na, bool a_scalar, const B* b, uint32_t nb, bool b_scalar) {
    const uint32_t n = a_scalar ? nb : na;
    ColRef<R> out = alloc_like_input<R>(a_scalar ? (const void*)b : (const void*)a, n);
    if (n) check(aqgpu_vec_binary(op, dtype_of<A>(), a, a_scalar, dtype_of<B>(), b, b_scalar, dtype_of<R>(), out.container, n));
    return out;
}
}  // namespace aq
#define AQ_VECOP(SYM, OP, VV, VS)                                                                                                   \
    template <class A, class B, class = std::enable_if_t<aq::is_vec_cell<A>::value && aq::is_vec_cell<B>::value && !std::is_same<A, B>::value>> \
    auto operator SYM(const ColRef<A>& a, const ColRef<B>& b) { using K [[maybe_unused]] = typename aq::coercion<aq::cell_t<A>, aq::cell_t<B>>::type; \
        return aq::vec_op<VV>(OP, a.container, a.size, false, b.container, b.size, false); }                                        \
    template <class A, class B, class = std::enable_if_t<aq::is_vec_cell<A>::value && std::is_arithmetic<B>::value>>                \
    auto operator SYM(const ColRef<A>& a, const B& b) { using K [[maybe_unused]] = typename aq::coercion<aq::cell_t<A>, aq::cell_t<B>>::type;        \
        const aq::cell_t<B> s = b; return aq::vec_op<VS>(OP, a.container, a.size, false, &s, 1u, true); }                           \
    template <class A, class B, class = std::enable_if_t<aq::is_vec_cell<A>::value && std::is_arithmetic<B>::value>>                \
    auto operator SYM(const B& b, const ColRef<A>& a) { using K [[maybe_unused]] = typename aq::coercion<aq::cell_t<A>, aq::cell_t<B>>::type;        \
        const aq::cell_t<B> s = b; return aq::vec_op<VS>(OP, &s, 1u, true, a.container, a.size, false); }
AQ_VECOP(+, AQ_V_ADD, K, K)
AQ_VECOP(-, AQ_V_SUB, K, K)
AQ_VECOP(*, AQ_V_MUL, K, types::GetLongType<K>)
AQ_VECOP(/, AQ_V_DIV, K, types::GetFPType<K>)
AQ_VECOP(>, AQ_V_GT, bool, bool) AQ_VECOP(<, AQ_V_LT, bool, bool) AQ_VECOP(>=, AQ_V_GTE, bool, bool)
AQ_VECOP(<=, AQ_V_LTE, bool, bool) AQ_VECOP(==, AQ_V_EQ, bool, bool) AQ_VECOP(!=, AQ_V_NEQ, bool, bool)
#undef AQ_VECOP
// same-type column pairs (operator- above keeps its aqgpu_vec_sub fast path)
#define AQ_VECOP_SAME(SYM, OP, RT)                                                                                                  \
    template <class T, class = std::enable_if_t<aq::is_vec_cell<T>::value>>                                                        \
    ColRef<RT> operator SYM(const ColRef<T>& a, const ColRef<T>& b) { return aq::vec_op<RT>(OP, a.container, a.size, false, b.container, b.size, false); }
AQ_VECOP_SAME(+, AQ_V_ADD, T) AQ_VECOP_SAME(*, AQ_V_MUL, T) AQ_VECOP_SAME(/, AQ_V_DIV, T)
AQ_VECOP_SAME(>, AQ_V_GT, bool) AQ_VECOP_SAME(<, AQ_V_LT, bool) AQ_VECOP_SAME(>=, AQ_V_GTE, bool)
AQ_VECOP_SAME(<=, AQ_V_LTE, bool) AQ_VECOP_SAME(==, AQ_V_EQ, bool) AQ_VECOP_SAME(!=, AQ_V_NEQ, bool)
#undef AQ_VECOP_SAME
// aqop_x(lhs, rhs, ret): the root of an emitted expression writes into a container of the emitter's type (table.h:949-973)
#define AQ_AQOP(NAME, OP)                                                                                                           \
    template <class A, class B, class R> void aqop_##NAME(const ColRef<A>& a, const ColRef<B>& b, ColRef<R>& ret) {                 \
        const int rt = sizeof(R) == 1 ? AQ_BOOL : aq::dtype_of<R>();          /* the emitter's ByteT for logical results: 0 / 1 cells */ \
        if (ret.size) aq::check(aqgpu_vec_binary(OP, aq::dtype_of<A>(), a.container, 0, aq::dtype_of<B>(), b.container, 0, rt, ret.container, ret.size)); }
AQ_AQOP(add, AQ_V_ADD) AQ_AQOP(minus, AQ_V_SUB) AQ_AQOP(mul, AQ_V_MUL) AQ_AQOP(div, AQ_V_DIV) AQ_AQOP(and, AQ_V_AND) AQ_AQOP(or, AQ_V_OR)
AQ_AQOP(xor, AQ_V_XOR) AQ_AQOP(gt, AQ_V_GT) AQ_AQOP(lt, AQ_V_LT) AQ_AQOP(gte, AQ_V_GTE) AQ_AQOP(lte, AQ_V_LTE) AQ_AQOP(eq, AQ_V_EQ) AQ_AQOP(neq, AQ_V_NEQ)
#undef AQ_AQOP
// sums over a comparison result, the session counter of tests/network.a:11 `sums(deltas(_time) > 120)`.  The running count is
// returned as int64 cells (the reference's GetLongType<bool> is a 128-bit integer) so that it can serve as a key column.
inline ColRef<long long> sums(const ColRef<bool>& c) {
    const int zero = 0;
    ColRef<int> wide = aq::vec_op<int>(AQ_V_ADD, c.container, c.size, false, &zero, 1u, true);
    auto g = aq::whole_column(wide);
    ColRef<long long> out = aq::alloc_like_input<long long>(c.container, c.size);
    if (c.size) aq::check(aqgpu_group_scan(g.handle->h, AQ_SUMS, 0, AQ_INT32, wide.container, AQ_INT64, out.container));
    return out;
}
template <class T, class = std::enable_if_t<(aq::dtype_of<T>() >= 0 && aq::dtype_of<T>() != AQ_INT128 && aq::dtype_of<T>() != AQ_BOOL && aq::dtype_of<T>() != AQ_UINT32)>>
ColRef<double> sqrt(const ColRef<T>& x) {                                                           // aggregations.h:34-47
    ColRef<double> out = aq::alloc_like_input<double>(x.container, x.size);
    if (x.size) aq::check(aqgpu_vec_unary(AQ_U_SQRT, aq::dtype_of<T>(), x.container, 0, out.container, x.size));
    return out;
}
template <class T, class = std::enable_if_t<(aq::dtype_of<T>() >= 0 && aq::dtype_of<T>() != AQ_INT128 && aq::dtype_of<T>() != AQ_BOOL && aq::dtype_of<T>() != AQ_UINT32)>>
ColRef<T> truncate(const ColRef<T>& x, const uint32_t precision) {                                  // aggregations.h:58-69
    ColRef<T> out = aq::alloc_like_input<T>(x.container, x.size);
    if (x.size) aq::check(aqgpu_vec_unary(AQ_U_TRUNCATE, aq::dtype_of<T>(), x.container, precision, out.container, x.size));
    return out;
}
inline ColRef<double> pow(const ColRef<double>& x, double p) {                                      // q9: pow(corr(v1, v2), 2)
    ColRef<double> out = aq::alloc_like_input<double>(x.container, x.size);
    if (x.size) aq::check(aqgpu_vec_pow(x.container, p, out.container, x.size));
    return out;
}
template <class T> ColRef<T> second(const aq::GroupedCol<T>& c) { return aq::reduce_col<T, T>(c, AQ_SECOND); }   // subvec(v, 0, 2)[1], h2o q8

namespace aq {
// Several vector-valued expressions of one group-by evaluated by ONE library call (aqgpu_group_window): what the
// whole-column emitter produces for `SELECT avgs(5, price), sums(5, quantity), ..., max(price - mins(price))`.
// add() / add_reduced() allocate the result column and queue the expression; run() evaluates all of them.
struct WindowBatch {
    std::shared_ptr<GroupsHandle> handle;
    std::vector<aqgpu_window_t> ops;
    template <class G> explicit WindowBatch(const G& g) : handle(g.handle) {}
    // scan op over col[g]; minus: the vector is col - scan(col)
    template <class R, class T> VecCol<R> add(int op, uint32_t w, const ColRef<T>& col, bool minus = false) {
        VecCol<R> out{alloc_like_input<R>(col.container, handle->nrows), handle};
        ops.push_back(aqgpu_window_t{op, w, dtype_of<T>(), col.container, minus ? 1 : 0, -1, dtype_of<R>(), out.flat.container});
        return out;
    }
    // max / min per group of the vector (aggregations.h:71-86 applied to the vector-valued expression)
    template <class T> ColRef<T> add_reduced(int op, uint32_t w, const ColRef<T>& col, bool minus, int reduce) {
        ColRef<T> out = alloc_like_input<T>(col.container, handle->ngroups);
        ops.push_back(aqgpu_window_t{op, w, dtype_of<T>(), col.container, minus ? 1 : 0, reduce, dtype_of<T>(), out.container});
        return out;
    }
    // ASSUMING ASC / DESC col (common/groupby.py:54-59 -> TableInfo::order_by per group): evaluated inside the window call
    const void* order_col = nullptr; int order_type = 0, order_desc = 0;
    template <class T> void assuming(const ColRef<T>& col, bool desc = false) { order_col = col.container; order_type = dtype_of<T>(); order_desc = desc ? 1 : 0; }
    void run() {
        if (!ops.empty() && handle->nrows) {
            if (order_col) check(aqgpu_group_window_ordered(handle->h, order_col, order_type, order_desc, ops.data(), (int)ops.size()));
            else check(aqgpu_group_window(handle->h, ops.data(), (int)ops.size()));
        }
        ops.clear();
    }
};
}  // namespace aq

namespace aq {
// row lists in ascending row order (a table the SQL side has sorted by time): AQ_ROWS_ASC
template <class... Ts> Groups<Ts...> group_by_rows_ascending(const ColRef<Ts>&... cols) {
    Groups<Ts...> g;
    g.handle = build_groups<Ts...>(std::index_sequence_for<Ts...>(), AQ_ROWS_ASC, cols...);
    g.size = g.handle->ngroups;
    return g;
}
// Cells of a result TABLE are host-resident (print / monetdb_append_table read them): a scalar column comes back
// as it is, a vector-valued one as vector_type cells over its flat CSR buffer (engine/ast.py:696-718).
template <class T> ColRef<T> table_cells(const ColRef<T>& c) {
    if (!c.on_device()) { ColRef<T> h; h.init_from(c.size, c.container); return h; }
    const std::vector<T> v = c.to_host();
    ColRef<T> h(c.size);
    if (c.size) memcpy((void*)h.container, (const void*)v.data(), sizeof(T) * c.size);
    return h;
}
template <class R> ColRef<vector_type<R>> table_cells(VecCol<R>& vc) {
    if (vc.flat.on_device()) { ColRef<R>* h = new ColRef<R>(table_cells(vc.flat)); vc.flat.init_from(h->size, h->container); }
    return vc.views();
}
// every column of a result table as raw cells in <prefix>_<i>.bin (vector cells: the flat buffer); test support
template <class C> void dump_column(ColRef<C>& col, const char* prefix, size_t i) {
    FILE* f = fopen((std::string(prefix) + "_" + std::to_string(i) + ".bin").c_str(), "wb");
    if (!f) return;
    if constexpr (is_vector_type_impl<C>::value) { for (uint32_t r = 0; r < col.size; ++r) fwrite(col.container[r].container, sizeof(*col.container[r].container), col.container[r].size, f); }
    else fwrite(col.container, sizeof(C), col.size, f);
    fclose(f);
}
template <class... Ts, size_t... I> void dump_table_impl(TableInfo<Ts...>& t, const char* prefix, std::index_sequence<I...>) {
    (dump_column(t.template get_col<(int)I>(), prefix, I), ...);
}
template <class... Ts> void dump_table(TableInfo<Ts...>& t, const char* prefix) { dump_table_impl(t, prefix, std::index_sequence_for<Ts...>()); }
}  // namespace aq

// ---- Context / DataSource: the objects the real server hands to c(cxt) -------------------------------------
// Layout-compatible restatements of server/libaquery.h:118-161 and server/DataSource_conn.h:27-54 for the
// default build of the reference (THREADING=1 from aquery_config.py:29, no __AQ_USE_THREADEDGC__): same
// member order, same virtual-function order, so a dll.so compiled against THESE headers can be loaded by the
// reference's unmodified server.so (server/server.cpp:296-307).  The offsets are pinned by the static_asserts
// below and compared with the reference's own headers in tests/test_host_layout.py.
#ifndef AQGPU_HOST_NO_THREADING
#define AQGPU_HOST_THREADING 1
#endif
enum Log_level : int { LOG_INFO, LOG_ERROR, LOG_SILENT };                                   // libaquery.h:60-64
enum Backend_Type : int { BACKEND_AQuery, BACKEND_MonetDB, BACKEND_MariaDB, BACKEND_DuckDB, BACKEND_SQLite, BACKEND_TOTAL };
struct QueryStats { long long monet_time, postproc_time; };                                 // libaquery.h:78-81
struct Config {                                                                             // libaquery.h:82-88
    int running, new_query, server_mode, backend_type, has_dll, n_buffers;
    QueryStats stats;
    int buffer_sizes[];
};
struct Session {                                                                            // libaquery.h:98-105
    struct Statistic { size_t total_active, cnt_object, total_alloc; } stats;
    void* memory_map;
};
struct StoredProcedure {                                                                    // libaquery.h:107-112
    uint32_t cnt, postproc_modules;
    char** queries;
    const char* name;
    void** __rt_loaded_modules;
};
struct Trigger;
struct IntervalBasedTriggerHost;
struct CallbackBasedTriggerHost;
struct uColRef;

struct Context;
struct DataSource {                                                                         // DataSource_conn.h:27-54
    void* server = nullptr;
    Context* cxt = nullptr;
    bool status = false;
    char* query = nullptr;
    Backend_Type DataSourceType = BACKEND_AQuery;
    void* res = nullptr;
    void* ret_col = nullptr;
    long long cnt = 0;
    const char* last_error = nullptr;
    void* handle = nullptr;
    DataSource() = default;
    // vtable order of the reference: connect, exec, getCol, getDSTable, close, haserror, print_results, destructor
    virtual void connect(Context* cxt) = 0;
    virtual void exec(const char* q) = 0;
    virtual void* getCol(int col_idx, int type) = 0;
    virtual void getDSTable(const char* name, void* tbl) = 0;
    virtual void close() = 0;
    virtual bool haserror() = 0;
    virtual void print_results(const char* = " ", const char* = "\n", uint32_t = 0xffffffffu) {}
    virtual ~DataSource() {}
};

struct Context {                                                                            // libaquery.h:118-161
    typedef int (*printf_type)(const char* format, ...);
    void* module_function_maps = nullptr;
    Config* cfg = nullptr;
    int n_buffers = 0, *sz_bufs = nullptr;
    void** buffers = nullptr;
    void* curr_server = nullptr;
    void* alt_server[BACKEND_TOTAL] = {nullptr};
    Log_level log_level = LOG_INFO;
    Session current{};
    const char* aquery_root_path = nullptr;
#ifdef AQGPU_HOST_THREADING
    void* thread_pool = nullptr;
#endif
    void* gc = nullptr;
    printf_type print = &printf;
    Context() {}
    virtual ~Context() {}                              // the only virtual of the reference's Context
    // maps of the reference (a libstdc++ unordered_map has one size whatever it maps to): tables, cols,
    // stored_proc, triggers.  The generated code reads `tables` (cached TableInfo<...>*, server.cpp:552-583).
    std::unordered_map<std::string, void*> tables;
    std::unordered_map<std::string, uColRef*> cols;
    std::unordered_map<std::string, StoredProcedure> stored_proc;
    std::unordered_map<std::string, void*> triggers;   // reference: unordered_map<string, Trigger>
    IntervalBasedTriggerHost* it_host = nullptr;
    CallbackBasedTriggerHost* ct_host = nullptr;
};
#if defined(__x86_64__) && defined(__GLIBCXX__)
#define AQ_OFFSET_IS(T, m, v) static_assert(__builtin_offsetof(T, m) == (v), "layout of " #T "::" #m " differs from the reference")
#pragma GCC diagnostic push
#pragma GCC diagnostic ignored "-Winvalid-offsetof"
AQ_OFFSET_IS(DataSource, server, 8); AQ_OFFSET_IS(DataSource, cxt, 16); AQ_OFFSET_IS(DataSource, status, 24);
AQ_OFFSET_IS(DataSource, query, 32); AQ_OFFSET_IS(DataSource, DataSourceType, 40); AQ_OFFSET_IS(DataSource, res, 48);
AQ_OFFSET_IS(DataSource, ret_col, 56); AQ_OFFSET_IS(DataSource, cnt, 64); AQ_OFFSET_IS(DataSource, last_error, 72);
AQ_OFFSET_IS(DataSource, handle, 80);
static_assert(sizeof(DataSource) == 88, "DataSource size differs from the reference");
AQ_OFFSET_IS(Context, module_function_maps, 8); AQ_OFFSET_IS(Context, cfg, 16); AQ_OFFSET_IS(Context, n_buffers, 24);
AQ_OFFSET_IS(Context, sz_bufs, 32); AQ_OFFSET_IS(Context, buffers, 40); AQ_OFFSET_IS(Context, curr_server, 48);
AQ_OFFSET_IS(Context, alt_server, 56); AQ_OFFSET_IS(Context, log_level, 96); AQ_OFFSET_IS(Context, current, 104);
AQ_OFFSET_IS(Context, aquery_root_path, 136);
#ifdef AQGPU_HOST_THREADING
AQ_OFFSET_IS(Context, thread_pool, 144); AQ_OFFSET_IS(Context, gc, 152); AQ_OFFSET_IS(Context, print, 160);
AQ_OFFSET_IS(Context, tables, 168); AQ_OFFSET_IS(Context, cols, 224); AQ_OFFSET_IS(Context, it_host, 392);

// ---- HBM column store behind cxt->tables (north_star item 1) ---------------------------------------------------------
// The GPU build's counterpart of the 'C' cache message (server/server.cpp:552-583: `server->getDSTable(name, tbl);
// cxt->tables[name] = tbl;`): every column of a host-resident table is ingested into the HBM column store ONCE and the
// device-backed TableInfo is registered under the table's name.  The emitter's cached-table route
// (engine/ast.py:344-365: `static_cast<TableInfo<...>*>(cxt->tables["name"])` + `->get_col<i>()`) then hands HBM-resident
// ColRefs to the post-processor; results live where their inputs live.
namespace aq {
template <class... Ts, size_t... I>
inline void cache_cols(TableInfo<Ts...>& dev, TableInfo<Ts...>& host, std::index_sequence<I...>) {
    ((void)[&] {
        auto d = host.template get_col<(int)I>().to_device();
        auto& c = dev.template get_col<(int)I>();
        c.init_from(d.size, d.container);              // the cached table keeps the column for the session
        c.name = host.template get_col<(int)I>().name;
        d.capacity = 0;
    }(), ...);
}
template <class... Ts> TableInfo<Ts...>* cache_table(Context* cxt, const char* name, TableInfo<Ts...>& host) {
    auto* dev = new TableInfo<Ts...>(name, nullptr);
    cache_cols(*dev, host, std::index_sequence_for<Ts...>());
    cxt->tables[name] = dev;
    return dev;
}
}  // namespace aq
static_assert(sizeof(Context) == 408, "Context size differs from the reference");
#endif
#pragma GCC diagnostic pop
#undef AQ_OFFSET_IS
#endif
