// plan.inl -- b200_plan_create / b200_plan_launch / b200_plan_destroy: a generated stencil as ONE C-ABI object.
//
// The counterpart of the reference's per-stencil pybind11 entry `run_computation(domain, field, origin, ..., exec_info)`
// (/root/reference/src/gt4py/cartesian/backend/gtc_common.py:79-101, call site :157-163): a host in any language hands
// over the plan descriptor the backend wrote at build time (`plan.desc`, text) plus the cubin, and then launches with
// `b200_field[]` / `b200_scalar[]` -- argument packing, entry-point choice (generic / unit-stride / TMA-staged), grid
// shapes, the scratch for the OIR temporaries, TMA descriptors and the spill areas of the fused column kernels are the
// plan's business (SURVEY.md 8(b) "Runtime call", "Ownership").
//
// Written against the C ABI only (b200_malloc, b200_module_load, b200_launch, b200_tmap_encode, ...), no CUDA header:
// csrc/plan.cu compiles it into libgt4py_b200.so, tests/emu_abi.cpp compiles the same text on top of its host doubles.
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <array>
#include <map>
#include <mutex>
#include <sstream>
#include <string>
#include <vector>

namespace b200plan {

static thread_local char g_perr[512];
static int fail(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_perr, sizeof(g_perr), fmt, ap);
    va_end(ap);
    B200_PLAN_SET_ERROR(g_perr);
    return code;
}

static int dtype_size(int dt)
{
    switch (dt) {
    case B200_BOOL: case B200_INT8: return 1;
    case B200_INT16: return 2;
    case B200_INT32: case B200_FLOAT32: return 4;
    case B200_INT64: case B200_FLOAT64: return 8;
    }
    return 0;
}

struct FieldSpec {
    std::string name;
    int dtype = 0, dims[3] = {0, 0, 0}, ndata = 0, ndim = 0;
};
struct ScalarSpec {
    std::string name;
    int dtype = 0;
};
struct TempSpec {
    std::string name;
    int dtype = 0, dims[3] = {0, 0, 0};
    std::vector<int> data;
    int li = 0, hi = 0, lj = 0, hj = 0;
};
enum MemberKind { M_FIELD, M_TMAP, M_ORIGIN, M_SCALAR, M_CF_SPILL, M_CF_SMEM, M_LEVEL };
struct Member {
    MemberKind kind;
    std::string name, ref, kernel;  // ref: field name (tmap / origin); kernel: owning kernel (tmap / cf_*)
    int nstr = 0, dtype = 0;
    uint32_t box[3] = {0, 0, 0};
    size_t offset = 0;              // byte offset in struct Args
    int field = -1, temp = -1, scalar = -1;  // resolved indices
};
enum KernelKind { K_POINT, K_COLUMN, K_TILE, K_COLFUSE, K_LEVEL, K_MARCH };
struct Kernel {
    std::string name;
    KernelKind kind;
    int ilo = 0, ihi = 0, jlo = 0, jhi = 0;
    int rows = 1, ti = 0, tj = 0, bx = 0, by = 0, smem = 0, has_tma = 0, smem_tma = 0, ring_slots = 0, stack_values = 0, stack_tmem = 128;
    void *entry = nullptr, *entry_u = nullptr, *entry_t = nullptr;
    void *spill = nullptr;
    size_t spill_bytes = 0;
    // `level` kernels: one stage of a sequential loop run level by level
    int wo = 0, warps = 0, strip = 0;  // `march` kernels: result columns per warp, warps per CTA, rows per strip
    int loop = -1, backward = 0;
    std::vector<std::array<int, 4>> intervals;  // (start level, offset, end level, offset); level 0 = START, 1 = END
    bool has_level(int64_t lvl, int64_t nk) const
    {
        for (const auto &v : intervals)
            if (nk * v[0] + v[1] <= lvl && lvl < nk * v[2] + v[3]) return true;
        return false;
    }
};
struct Scratch {
    std::vector<void *> base;      // allocations
    std::vector<void *> origin;    // pointer of the temporaries' compute origin
    std::vector<std::vector<int64_t>> strides;
};
struct TmapSlot {
    const void *ptr;
    int64_t shape[3], stride[3];
    int member;
    void *dev;
    int shift;
    bool ok;
};
struct Library {
    std::string kernel;
    int limiter = 0, min_k = 1;
    std::map<std::string, std::string> roles;  // role -> API argument name
};

struct Plan {
    std::string name;
    std::vector<FieldSpec> fields;
    std::vector<ScalarSpec> scalars;
    std::vector<TempSpec> temps;
    std::vector<Member> members;
    std::vector<Kernel> kernels;
    bool has_library = false, use_library = true, no_tma = false;
    Library lib;
    size_t args_bytes = 0, off_ni = 0;
    std::vector<unsigned char> args;
    void *module = nullptr;
    std::map<std::vector<int64_t>, Scratch> scratch;  // keyed by (ni, nj, nk, stream)
    std::vector<TmapSlot> tmaps;
    std::string last_entries;
    int sm_count = 0;
    std::mutex lock;
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// struct Args as nvcc lays it out: members in order, each at its natural alignment, int32 ni, nj, nk last
static void layout(Plan &p)
{
    size_t off = 0, maxal = 8;
    for (Member &m : p.members) {
        size_t sz = 0, al = 8;
        switch (m.kind) {
        case M_FIELD: {  // T *name; int64_t name_s[nstr];
            off = align_up(off, 8);
            m.offset = off;
            off += 8 + 8 * (size_t)m.nstr;
            continue;
        }
        case M_TMAP: case M_CF_SPILL: sz = 8; al = 8; break;
        case M_ORIGIN: sz = 12; al = 4; break;
        case M_CF_SMEM: case M_LEVEL: sz = 4; al = 4; break;
        case M_SCALAR: sz = al = (size_t)dtype_size(m.dtype); break;
        }
        off = align_up(off, al);
        m.offset = off;
        off += sz;
    }
    off = align_up(off, 4);
    p.off_ni = off;
    off += 12;
    p.args_bytes = align_up(off, maxal);
    p.args.assign(p.args_bytes, 0);
}

static int find_kernel(Plan &p, const std::string &name)
{
    for (size_t n = 0; n < p.kernels.size(); ++n)
        if (p.kernels[n].name == name) return (int)n;
    return -1;
}

static int parse(Plan &p, const std::string &text)
{
    std::istringstream in(text);
    std::string line, tok;
    bool header = false, ended = false;
    while (std::getline(in, line)) {
        std::istringstream ls(line);
        if (!(ls >> tok) || tok[0] == '#') continue;
        if (tok == "b200plan") {
            int version = 0;
            ls >> version >> p.name;
            if (version != 1) return fail(B200_EINVAL, "plan descriptor: unknown version %d", version);
            header = true;
        } else if (tok == "field") {
            FieldSpec f;
            ls >> f.name >> f.dtype >> f.dims[0] >> f.dims[1] >> f.dims[2] >> f.ndata;
            f.ndim = f.dims[0] + f.dims[1] + f.dims[2] + f.ndata;
            if (!ls || !dtype_size(f.dtype)) return fail(B200_EINVAL, "plan descriptor: bad field line '%s'", line.c_str());
            p.fields.push_back(f);
        } else if (tok == "scalar") {
            ScalarSpec s;
            ls >> s.name >> s.dtype;
            if (!ls || !dtype_size(s.dtype)) return fail(B200_EINVAL, "plan descriptor: bad scalar line '%s'", line.c_str());
            p.scalars.push_back(s);
        } else if (tok == "temp") {
            TempSpec t;
            int nd = 0;
            ls >> t.name >> t.dtype >> t.dims[0] >> t.dims[1] >> t.dims[2] >> nd;
            for (int n = 0; n < nd; ++n) { int d; ls >> d; t.data.push_back(d); }
            ls >> t.li >> t.hi >> t.lj >> t.hj;
            if (!ls || !dtype_size(t.dtype)) return fail(B200_EINVAL, "plan descriptor: bad temp line '%s'", line.c_str());
            p.temps.push_back(t);
        } else if (tok == "member") {
            Member m;
            std::string kind;
            ls >> kind >> m.name;
            if (kind == "field") { m.kind = M_FIELD; ls >> m.nstr; }
            else if (kind == "tmap") { m.kind = M_TMAP; ls >> m.ref >> m.kernel >> m.box[0] >> m.box[1] >> m.box[2]; }
            else if (kind == "origin") { m.kind = M_ORIGIN; ls >> m.ref; }
            else if (kind == "scalar") { m.kind = M_SCALAR; ls >> m.dtype; }
            else if (kind == "cf_spill") { m.kind = M_CF_SPILL; ls >> m.kernel; }
            else if (kind == "cf_smem") { m.kind = M_CF_SMEM; ls >> m.kernel; }
            else if (kind == "level") { m.kind = M_LEVEL; }
            else return fail(B200_EINVAL, "plan descriptor: unknown member kind '%s'", kind.c_str());
            if (!ls) return fail(B200_EINVAL, "plan descriptor: bad member line '%s'", line.c_str());
            p.members.push_back(m);
        } else if (tok == "kernel") {
            Kernel k;
            std::string kind;
            ls >> k.name >> kind >> k.ilo >> k.ihi >> k.jlo >> k.jhi;
            if (kind == "point") { k.kind = K_POINT; ls >> k.rows; }
            else if (kind == "column") { k.kind = K_COLUMN; }
            else if (kind == "march") { k.kind = K_MARCH; ls >> k.wo >> k.warps >> k.strip >> k.smem; }
            else if (kind == "level") {
                int n = 0;
                k.kind = K_LEVEL;
                ls >> k.rows >> k.loop >> k.backward >> n;
                for (int v = 0; v < n; ++v) {
                    std::array<int, 4> b{};
                    ls >> b[0] >> b[1] >> b[2] >> b[3];
                    k.intervals.push_back(b);
                }
            }
            else if (kind == "tile") { k.kind = K_TILE; ls >> k.ti >> k.tj >> k.bx >> k.by >> k.smem >> k.has_tma >> k.smem_tma; }
            else if (kind == "colfuse") { k.kind = K_COLFUSE; ls >> k.bx >> k.by >> k.ring_slots >> k.stack_values >> k.stack_tmem >> k.has_tma >> k.smem_tma; }
            else return fail(B200_EINVAL, "plan descriptor: unknown kernel kind '%s'", kind.c_str());
            if (!ls) return fail(B200_EINVAL, "plan descriptor: bad kernel line '%s'", line.c_str());
            p.kernels.push_back(k);
        } else if (tok == "library") {
            int n = 0;
            ls >> p.lib.kernel >> p.lib.limiter >> p.lib.min_k >> n;
            for (int r = 0; r < n; ++r) {
                std::string kv;
                ls >> kv;
                size_t eq = kv.find('=');
                if (eq == std::string::npos) return fail(B200_EINVAL, "plan descriptor: bad library role '%s'", kv.c_str());
                p.lib.roles[kv.substr(0, eq)] = kv.substr(eq + 1);
            }
            p.has_library = true;
        } else if (tok == "end") {
            ended = true;
        } else {
            return fail(B200_EINVAL, "plan descriptor: unknown line '%s'", line.c_str());
        }
    }
    if (!header || !ended) return fail(B200_EINVAL, "plan descriptor: missing 'b200plan' header or 'end' line (truncated?)");
    // resolve names
    for (Member &m : p.members) {
        auto field_index = [&](const std::string &nm) {
            for (size_t n = 0; n < p.fields.size(); ++n) if (p.fields[n].name == nm) return (int)n;
            return -1;
        };
        if (m.kind == M_FIELD) {
            m.field = field_index(m.name);
            if (m.field < 0)
                for (size_t n = 0; n < p.temps.size(); ++n) if (p.temps[n].name == m.name) m.temp = (int)n;
            if (m.field < 0 && m.temp < 0) return fail(B200_EINVAL, "plan descriptor: member '%s' is neither field nor temporary", m.name.c_str());
        } else if (m.kind == M_TMAP || m.kind == M_ORIGIN) {
            m.field = field_index(m.ref);
            if (m.field < 0) return fail(B200_EINVAL, "plan descriptor: member '%s' refers to unknown field '%s'", m.name.c_str(), m.ref.c_str());
        } else if (m.kind == M_SCALAR) {
            for (size_t n = 0; n < p.scalars.size(); ++n) if (p.scalars[n].name == m.name) m.scalar = (int)n;
            if (m.scalar < 0) return fail(B200_EINVAL, "plan descriptor: scalar member '%s' not declared", m.name.c_str());
        }
        if ((m.kind == M_TMAP || m.kind == M_CF_SPILL || m.kind == M_CF_SMEM) && find_kernel(p, m.kernel) < 0)
            return fail(B200_EINVAL, "plan descriptor: member '%s' refers to unknown kernel '%s'", m.name.c_str(), m.kernel.c_str());
    }
    layout(p);
    return B200_OK;
}

}  // namespace b200plan

// ---- the three entry points -------------------------------------------------------------------------------------
namespace b200plan {

template <typename T> static void put(Plan &p, size_t off, T v) { memcpy(p.args.data() + off, &v, sizeof(T)); }

static int make_scratch(Plan &p, const int64_t dom[3], Scratch &s)
{
    for (const TempSpec &t : p.temps) {
        const int item = dtype_size(t.dtype);
        // I-contiguous (layout_map (2, 1, 0)), data dimensions outermost; the I extent padded to whole 128-byte lines
        // and the compute origin placed on a line boundary, like gt4py_b200.storage.empty(aligned_index=(li, lj, 0))
        const int64_t ei = t.dims[0] ? dom[0] + t.li + t.hi : 1, ej = t.dims[1] ? dom[1] + t.lj + t.hj : 1;
        const int64_t ek = t.dims[2] ? dom[2] : 1;
        const int64_t per = 128 / item, pi = (ei + per - 1) / per * per;
        int64_t nd = 1;
        for (int d : t.data) nd *= d;
        const size_t bytes = (size_t)(pi * ej * ek * nd) * item + 256;
        void *base = nullptr;
        int rc = b200_malloc(&base, bytes);
        if (rc) return rc;
        const int64_t lead = t.dims[0] ? t.li : 0;
        const size_t shift = (size_t)((per - lead % per) % per) * item;  // so that element `li` of a row is line-aligned
        char *first = (char *)base + shift;
        std::vector<int64_t> st = {t.dims[0] ? 1 : 0, t.dims[1] ? pi : 0, t.dims[2] ? pi * ej : 0};
        int64_t acc = pi * ej * ek;
        std::vector<int64_t> ds(t.data.size());
        for (int n = (int)t.data.size() - 1; n >= 0; --n) { ds[n] = acc; acc *= t.data[n]; }
        st.insert(st.end(), ds.begin(), ds.end());
        s.base.push_back(base);
        s.origin.push_back(first + ((t.dims[0] ? t.li : 0) + (t.dims[1] ? t.lj * pi : 0)) * item);
        s.strides.push_back(st);
    }
    return B200_OK;
}

static void free_scratch(Scratch &s)
{
    for (void *b : s.base) b200_free(b);
    s.base.clear();
}

static int create(const char *desc, size_t desc_bytes, const void *image, size_t image_bytes, int flags, Plan **out)
{
    if (!desc || !image || !out) return fail(B200_EINVAL, "plan_create: null argument");
    Plan *p = new Plan();
    p->use_library = (flags & 1) == 0;
    p->no_tma = (flags & 2) != 0;
    int rc = parse(*p, std::string(desc, desc_bytes));
    if (!rc) rc = b200_module_load(image, image_bytes, &p->module);
    for (size_t n = 0; !rc && n < p->kernels.size(); ++n) {
        Kernel &k = p->kernels[n];
        rc = b200_module_get_kernel(p->module, k.name.c_str(), &k.entry);
        if (!rc) rc = b200_module_get_kernel(p->module, (k.name + "_u").c_str(), &k.entry_u);
        if (!rc && k.has_tma) rc = b200_module_get_kernel(p->module, (k.name + "_t").c_str(), &k.entry_t);
        if (rc) break;
        if (k.kind == K_COLFUSE) {
            rc = b200_kernel_set_smem(k.entry, B200_PLAN_COLFUSE_SMEM);
            if (!rc) rc = b200_kernel_set_smem(k.entry_u, B200_PLAN_COLFUSE_SMEM);
            if (!rc && k.has_tma) rc = b200_kernel_set_smem(k.entry_t, B200_PLAN_COLFUSE_SMEM);
        } else if (k.kind == K_TILE || k.kind == K_MARCH) {
            if (k.smem > 48 * 1024) {
                rc = b200_kernel_set_smem(k.entry, k.smem);
                if (!rc) rc = b200_kernel_set_smem(k.entry_u, k.smem);
            }
            if (!rc && k.has_tma && k.smem_tma > 48 * 1024) rc = b200_kernel_set_smem(k.entry_t, k.smem_tma);
        }
    }
    if (!rc) {
        int64_t props[8] = {0};
        rc = b200_device_props(props);
        p->sm_count = (int)props[0];
    }
    if (rc) {
        if (p->module) b200_module_unload(p->module);
        delete p;
        return rc;
    }
    *out = p;
    return B200_OK;
}

static void destroy(Plan *p)
{
    if (!p) return;
    for (auto &kv : p->scratch) free_scratch(kv.second);
    for (TmapSlot &t : p->tmaps) if (t.dev) b200_free(t.dev);
    for (Kernel &k : p->kernels) if (k.spill) b200_free(k.spill);
    if (p->module) b200_module_unload(p->module);
    delete p;
}

static int launch_library(Plan &p, const int64_t dom[3], const b200_field *fields, const b200_scalar *scalars, void *stream)
{
    auto field_of = [&](const char *role) -> const b200_field * {
        auto it = p.lib.roles.find(role);
        if (it == p.lib.roles.end()) return nullptr;
        for (size_t n = 0; n < p.fields.size(); ++n) if (p.fields[n].name == it->second) return &fields[n];
        return nullptr;
    };
    if (p.lib.kernel == "hdiff")
        return b200_hdiff(field_of("in_field"), field_of("out_field"), field_of("coeff"), dom, p.lib.limiter, 0, stream);
    if (p.lib.kernel == "vadv") {
        double dtr = 0.0;
        auto it = p.lib.roles.find("dtr_stage");
        for (size_t n = 0; it != p.lib.roles.end() && n < p.scalars.size(); ++n)
            if (p.scalars[n].name == it->second) dtr = scalars[n].value.f64;
        return b200_vadv(field_of("utens_stage"), field_of("u_stage"), field_of("wcon"), field_of("u_pos"), field_of("utens"),
                         dtr, dom, 0, stream);
    }
    if (p.lib.kernel == "tridiag")
        return b200_tridiag(field_of("inf"), field_of("diag"), field_of("sup"), field_of("rhs"), field_of("out"), dom, 0, stream);
    return fail(B200_EINVAL, "plan names an unknown library kernel '%s'", p.lib.kernel.c_str());
}

static int launch(Plan &p, const int64_t dom[3], const b200_field *fields, int nfields, const b200_scalar *scalars,
                  int nscalars, void *stream, int sync)
{
    std::lock_guard<std::mutex> guard(p.lock);
    if (!dom) return fail(B200_EINVAL, "plan_launch: domain is null");
    if (nfields != (int)p.fields.size() || nscalars != (int)p.scalars.size())
        return fail(B200_EINVAL, "plan_launch(%s): expected %zu fields and %zu scalars, got %d and %d", p.name.c_str(),
                    p.fields.size(), p.scalars.size(), nfields, nscalars);
    if ((nfields && !fields) || (nscalars && !scalars)) return fail(B200_EINVAL, "plan_launch: null argument array");
    const int64_t ni = dom[0], nj = dom[1], nk = dom[2];
    if (ni < 0 || nj < 0 || nk < 0 || ni >= (1ll << 31) || nj >= (1ll << 31) || nk >= (1ll << 31))
        return fail(B200_EINVAL, "plan_launch: bad domain %lld x %lld x %lld", (long long)ni, (long long)nj, (long long)nk);
    for (int n = 0; n < nfields; ++n) {
        const b200_field &f = fields[n];
        if (!f.ptr) continue;  // an unreferenced / optional field passed as None
        const FieldSpec &spec = p.fields[n];
        if (f.dtype != spec.dtype)
            return fail(B200_EINVAL, "plan_launch(%s): field '%s' has dtype code %d, the stencil expects %d", p.name.c_str(),
                        spec.name.c_str(), f.dtype, spec.dtype);
        if (f.ndim != spec.ndim)
            return fail(B200_EINVAL, "plan_launch(%s): field '%s' has %d dimensions, the stencil expects %d", p.name.c_str(),
                        spec.name.c_str(), f.ndim, spec.ndim);
        // extents against domain + halo are checked by the caller's validation, as in the reference (stencil_object.py:
        // 345-497 runs before `run`): the plan does not know the halos, and the reference accepts origins that only the
        // stencil's own offsets bring back inside the array (test_code_generation.py TestNegativeOrigin)
    }
    p.last_entries.clear();
    if (p.has_library && p.use_library && nk >= p.lib.min_k) {
        int rc = launch_library(p, dom, fields, scalars, stream);
        if (!rc) p.last_entries = std::string("library:") + b200_last_kernel();
        if (!rc && sync) rc = stream ? b200_stream_sync(stream) : b200_device_sync();
        return rc;
    }
    // ---- scratch of the OIR temporaries for this (domain, stream) -------------------------------------------------
    Scratch *scr = nullptr;
    if (!p.temps.empty()) {
        std::vector<int64_t> key = {ni, nj, nk, (int64_t)(intptr_t)stream};
        auto it = p.scratch.find(key);
        if (it == p.scratch.end()) {
            if (p.scratch.size() >= 8) {  // bounded: drop the oldest set
                free_scratch(p.scratch.begin()->second);
                p.scratch.erase(p.scratch.begin());
            }
            Scratch s;
            int rc = make_scratch(p, dom, s);
            if (rc) { free_scratch(s); return rc; }
            it = p.scratch.emplace(key, std::move(s)).first;
        }
        scr = &it->second;
    }
    // ---- argument block ----------------------------------------------------------------------------------------------
    bool unit = true;  // every field I-contiguous and < 2^31 elements -> the <kernel>_u entry points
    for (Member &m : p.members) {
        if (m.kind == M_FIELD && m.field >= 0) {
            const b200_field &f = fields[m.field];
            const FieldSpec &spec = p.fields[m.field];
            if (!f.ptr) { put<void *>(p, m.offset, nullptr); continue; }
            const int item = dtype_size(spec.dtype);
            int64_t strides[8] = {0};
            char *ptr = (char *)f.ptr;
            int pos = 0;
            for (int d = 0; d < 3; ++d)
                if (spec.dims[d]) {
                    strides[d] = f.stride[pos];
                    ptr += (int64_t)f.origin[pos] * f.stride[pos] * item;
                    ++pos;
                }
            for (int n = 0; n < spec.ndata; ++n) strides[3 + n] = f.stride[pos + n];
            int64_t span = 0;
            for (int n = 0; n < f.ndim; ++n) span += (f.shape[n] - 1) * (f.stride[n] < 0 ? -f.stride[n] : f.stride[n]);
            unit = unit && (!spec.dims[0] || strides[0] == 1) && span < (1ll << 31) - (1ll << 20);
            put<void *>(p, m.offset, ptr);
            memcpy(p.args.data() + m.offset + 8, strides, 8 * (size_t)m.nstr);
        } else if (m.kind == M_FIELD) {
            put<void *>(p, m.offset, scr->origin[m.temp]);
            memcpy(p.args.data() + m.offset + 8, scr->strides[m.temp].data(), 8 * (size_t)m.nstr);
            unit = unit && (scr->strides[m.temp][0] == 0 || scr->strides[m.temp][0] == 1);
        } else if (m.kind == M_SCALAR) {
            memcpy(p.args.data() + m.offset, &scalars[m.scalar].value, (size_t)dtype_size(m.dtype));
        }
    }
    put<int32_t>(p, p.off_ni, (int32_t)ni);
    put<int32_t>(p, p.off_ni + 4, (int32_t)nj);
    put<int32_t>(p, p.off_ni + 8, (int32_t)nk);
    // ---- TMA descriptors of the staged inputs (cached per field layout; an incompatible layout -> plain entry) --------
    std::vector<int> tma_ok(p.kernels.size(), -1);  // -1 no staged input, 0 not usable, 1 usable
    if (unit && !p.no_tma) {
        std::map<int, int> shifts;
        for (size_t mi = 0; mi < p.members.size(); ++mi) {
            Member &m = p.members[mi];
            if (m.kind != M_TMAP) continue;
            const b200_field &f = fields[m.field];
            const int kn = find_kernel(p, m.kernel);
            TmapSlot *hit = nullptr;
            for (TmapSlot &t : p.tmaps)
                if (t.member == (int)mi && t.ptr == f.ptr && !memcmp(t.shape, f.shape, 24) && !memcmp(t.stride, f.stride, 24)) hit = &t;
            if (!hit) {
                TmapSlot t;
                t.ptr = f.ptr; t.member = (int)mi; t.dev = nullptr; t.shift = 0; t.ok = false;
                memcpy(t.shape, f.shape, 24);
                memcpy(t.stride, f.stride, 24);
                alignas(64) unsigned char blob[128];
                int32_t shift = 0;
                int rc = f.ptr ? b200_tmap_encode(&f, m.box, blob, &shift) : B200_EUNSUP;
                if (rc == B200_OK) {
                    rc = b200_malloc(&t.dev, 128);
                    if (!rc) rc = b200_memcpy_h2d(t.dev, blob, 128, nullptr);
                    if (rc) return rc;
                    t.shift = shift;
                    t.ok = true;
                } else if (rc != B200_EUNSUP) {
                    return rc;
                }
                if (p.tmaps.size() >= 256) {  // bounded: drop the oldest descriptor
                    if (p.tmaps.front().dev) b200_free(p.tmaps.front().dev);
                    p.tmaps.erase(p.tmaps.begin());
                }
                p.tmaps.push_back(t);
                hit = &p.tmaps.back();
            }
            if (!hit->ok) { tma_ok[kn] = 0; continue; }
            put<void *>(p, m.offset, hit->dev);
            shifts[m.field] = hit->shift;
            if (tma_ok[kn] < 0) tma_ok[kn] = 1;
        }
        for (Member &m : p.members)
            if (m.kind == M_ORIGIN) {
                const b200_field &f = fields[m.field];
                int32_t org[3] = {f.origin[0], f.origin[1], f.origin[2]};
                auto it = shifts.find(m.field);
                if (it != shifts.end()) org[0] += it->second;
                memcpy(p.args.data() + m.offset, org, 12);
            }
    }
    // ---- launches ---------------------------------------------------------------------------------------------------------
    void *argp[1] = {p.args.data()};
    for (size_t n = 0; n < p.kernels.size(); ++n) {
        Kernel &k = p.kernels[n];
        const int64_t wi = ni + k.ihi - k.ilo, wj = nj + k.jhi - k.jlo;
        if (k.kind == K_LEVEL) {
            // a sequential loop whose stages exchange values horizontally: `for k: for stage: all ij` -- the stages are
            // the consecutive `level` kernels of one loop id, each launched on one level (member `lvl`) at a time
            size_t end = n;
            while (end < p.kernels.size() && p.kernels[end].kind == K_LEVEL && p.kernels[end].loop == k.loop) ++end;
            size_t off_lvl = 0;
            for (Member &m : p.members) if (m.kind == M_LEVEL) off_lvl = m.offset;
            std::vector<char> noted(end - n, 0);
            for (int64_t t = 0; t < nk; ++t) {
                const int64_t lvl = k.backward ? nk - 1 - t : t;
                for (size_t g = n; g < end; ++g) {
                    Kernel &s = p.kernels[g];
                    const int64_t gi = ni + s.ihi - s.ilo, gj = nj + s.jhi - s.jlo, rows = 4 * (int64_t)s.rows;
                    if (gi <= 0 || gj <= 0 || !s.has_level(lvl, nk)) continue;
                    uint32_t grid[3] = {(uint32_t)((gi + 63) / 64), (uint32_t)((gj + rows - 1) / rows), 1}, block[3] = {64, 4, 1};
                    put<int32_t>(p, off_lvl, (int32_t)lvl);
                    if (!noted[g - n]) {
                        noted[g - n] = 1;
                        if (!p.last_entries.empty()) p.last_entries += ",";
                        p.last_entries += s.name + (unit ? "_u" : "");
                    }
                    int rc = b200_launch(unit ? s.entry_u : s.entry, grid, block, 0, stream, argp, 1);
                    if (rc) return rc;
                }
            }
            n = end - 1;
            continue;
        }
        if (wi <= 0 || wj <= 0) continue;
        uint32_t grid[3], block[3], smem = 0;
        void *entry = unit ? k.entry_u : k.entry;
        const char *sfx = unit ? "_u" : "";
        if (k.kind == K_POINT) {
            const int64_t rows = 4 * (int64_t)k.rows;
            block[0] = 64; block[1] = 4; block[2] = 1;
            grid[0] = (uint32_t)((wi + 63) / 64); grid[1] = (uint32_t)((wj + rows - 1) / rows); grid[2] = (uint32_t)nk;
        } else if (k.kind == K_MARCH) {
            const int64_t per_cta = (int64_t)k.wo * k.warps;
            if (per_cta <= 0 || k.strip <= 0) return fail(B200_EINVAL, "plan_launch: bad march geometry of kernel '%s'", k.name.c_str());
            block[0] = 32u * (uint32_t)k.warps; block[1] = 1; block[2] = 1;
            grid[0] = (uint32_t)((ni + per_cta - 1) / per_cta); grid[1] = (uint32_t)((nj + k.strip - 1) / k.strip); grid[2] = (uint32_t)nk;
            smem = (uint32_t)k.smem;  // the lanes' input rings
        } else if (k.kind == K_TILE) {
            block[0] = (uint32_t)k.bx; block[1] = (uint32_t)k.by; block[2] = 1;
            grid[0] = (uint32_t)((ni + k.ti - 1) / k.ti); grid[1] = (uint32_t)((nj + k.tj - 1) / k.tj); grid[2] = (uint32_t)nk;
            smem = (uint32_t)k.smem;
            if (tma_ok[n] == 1 && k.entry_t) { entry = k.entry_t; sfx = "_t"; smem = (uint32_t)k.smem_tma; }
        } else if (k.kind == K_COLFUSE) {
            // persistent fused FORWARD / BACKWARD kernel: a CTA per SM walks column tiles; the value stack beyond the
            // thread's share of Tensor Memory goes to shared memory, then to an L2-resident spill area
            const int64_t ntiles = ((wi + k.bx - 1) / k.bx) * ((wj + k.by - 1) / k.by);
            const int64_t ctas = ntiles < p.sm_count ? (ntiles > 0 ? ntiles : 1) : p.sm_count;
            const int64_t slot = (int64_t)k.bx * k.by * 8;  // one value (or one ring level) of every thread of the CTA
            int64_t need = nk * k.stack_values - k.stack_tmem;
            if (need < 0) need = 0;
            int64_t ring = (int64_t)k.ring_slots * slot;
            if (tma_ok[n] == 1 && k.entry_t) { entry = k.entry_t; sfx = "_t"; ring = k.smem_tma; }  // rings fed by TMA
            int64_t n_smem = (B200_PLAN_COLFUSE_SMEM - ring) / slot;
            if (n_smem > need) n_smem = need;
            if (n_smem < 0) n_smem = 0;
            smem = (uint32_t)(ring + n_smem * slot);
            size_t spill = (size_t)((need - n_smem) * ctas * k.bx * k.by * 8);
            if (spill < 256) spill = 256;
            if (spill > k.spill_bytes) {
                if (k.spill) b200_free(k.spill);
                k.spill = nullptr;
                k.spill_bytes = 0;
                int rc = b200_malloc(&k.spill, spill);
                if (rc) return rc;
                k.spill_bytes = spill;
            }
            for (Member &m : p.members) {
                if (m.kernel != k.name) continue;
                if (m.kind == M_CF_SPILL) put<void *>(p, m.offset, k.spill);
                if (m.kind == M_CF_SMEM) put<int32_t>(p, m.offset, (int32_t)n_smem);
            }
            // (the <k>_t entry runs with one more row of threads: the producer of its TMA rings, lowering_colfuse._emit_tma)
            block[0] = (uint32_t)k.bx; block[1] = (uint32_t)(k.by + (entry == k.entry_t && k.entry_t ? 1 : 0)); block[2] = 1;
            grid[0] = (uint32_t)ctas; grid[1] = 1; grid[2] = 1;
        } else {
            block[0] = 32; block[1] = 4; block[2] = 1;
            grid[0] = (uint32_t)((wi + 31) / 32); grid[1] = (uint32_t)((wj + 3) / 4); grid[2] = 1;
        }
        if (nk <= 0 && k.kind != K_COLUMN && k.kind != K_COLFUSE) continue;
        if (!p.last_entries.empty()) p.last_entries += ",";
        p.last_entries += k.name + sfx;
        int rc = b200_launch(entry, grid, block, smem, stream, argp, 1);
        if (rc) return rc;
    }
    if (sync) return stream ? b200_stream_sync(stream) : b200_device_sync();
    return B200_OK;
}

}  // namespace b200plan

extern "C" {
int b200_plan_create(const char *desc, size_t desc_bytes, const void *image, size_t image_bytes, int flags, void **plan)
{
    b200plan::Plan *p = nullptr;
    int rc = b200plan::create(desc, desc_bytes, image, image_bytes, flags, &p);
    if (!rc) *plan = p;
    return rc;
}
int b200_plan_launch(void *plan, const int64_t domain[3], const b200_field *fields, int nfields,
                     const b200_scalar *scalars, int nscalars, void *stream, int sync)
{
    if (!plan) return b200plan::fail(B200_EINVAL, "plan_launch: plan is null");
    return b200plan::launch(*(b200plan::Plan *)plan, domain, fields, nfields, scalars, nscalars, stream, sync);
}
int b200_plan_destroy(void *plan)
{
    b200plan::destroy((b200plan::Plan *)plan);
    return B200_OK;
}
int b200_plan_last_entries(void *plan, char *buf, size_t n)
{
    if (!plan || !buf || !n) return b200plan::fail(B200_EINVAL, "plan_last_entries: bad arguments");
    const std::string &s = ((b200plan::Plan *)plan)->last_entries;
    snprintf(buf, n, "%s", s.c_str());
    return B200_OK;
}
}
