// dx_flow.cu — flow artefact loading, statement planning/execution and route() bookkeeping.
// Reference call sites are listed in dx_flow.h; each function below names the lines it restates.
#include "dx_flow.h"
#include <chrono>
#include "dx_rules.h"
#include "dx_join.h"
#include <cmath>
#include <sstream>

using namespace dxsql;

// =======================================================================================
// artefact loaders
// =======================================================================================
static std::string strip_bom(const std::string& s) {
    if (s.size() >= 3 && (unsigned char)s[0] == 0xEF && (unsigned char)s[1] == 0xBB && (unsigned char)s[2] == 0xBF) return s.substr(3);
    return s;
}
static std::string trim(const std::string& s) {
    size_t a = 0, b = s.size();
    while (a < b && isspace((unsigned char)s[a])) a++;
    while (b > a && isspace((unsigned char)s[b - 1])) b--;
    return s.substr(a, b - a);
}
static std::vector<std::string> split_lines(const std::string& text) {
    std::vector<std::string> out; std::string cur;
    for (char c : text) { if (c == '\n') { if (!cur.empty() && cur.back() == '\r') cur.pop_back(); out.push_back(cur); cur.clear(); } else cur.push_back(c); }
    if (!cur.empty() && cur.back() == '\r') cur.pop_back();
    out.push_back(cur);
    return out;
}

// ConfigManager.loadConfig: "key=value" lines, '#' comments (reference: datax/config/ConfigManager.scala:98-135)
static void parse_conf(const std::string& text, std::map<std::string, std::string>* out) {
    for (auto& raw : split_lines(strip_bom(text))) {
        std::string line = trim(raw);
        if (line.empty() || line[0] == '#') continue;
        size_t eq = line.find('=');
        if (eq == std::string::npos || eq == 0) continue;
        (*out)[trim(line.substr(0, eq))] = trim(line.substr(eq + 1));
    }
}

// scala.concurrent.duration.Duration.create("5 minutes" | "60 second" | "1 hour" ...), as used by
// TimeWindowHandler.scala:33-35,53 and SettingDictionary.scala:45-46.  Returns whole seconds
// (route() only ever uses .toSeconds, CommonProcessorFactory.scala:161-163).
long long dx_parse_duration_seconds(const std::string& s0, bool* ok) {
    std::string s = trim(s0); *ok = false;
    size_t i = 0; while (i < s.size() && (isdigit((unsigned char)s[i]) || s[i] == '.')) i++;
    if (i == 0) return 0;
    double v = atof(s.substr(0, i).c_str());
    std::string u = lower(trim(s.substr(i)));
    double mul;
    if (u == "d" || u == "day" || u == "days") mul = 86400;
    else if (u == "h" || u == "hr" || u == "hrs" || u == "hour" || u == "hours") mul = 3600;
    else if (u == "m" || u == "min" || u == "mins" || u == "minute" || u == "minutes") mul = 60;
    else if (u == "s" || u == "sec" || u == "secs" || u == "second" || u == "seconds") mul = 1;
    else if (u == "ms" || u == "milli" || u == "millis" || u == "millisecond" || u == "milliseconds") mul = 1e-3;
    else return 0;
    *ok = true;
    return (long long)std::floor(v * mul);
}

// TransformSQLParser.parse (reference: datax/sql/TransformSqlParser.scala:27-95), line by line.
static bool word_boundary_contains(const std::string& text, const std::string& word) {
    auto isw = [](char c) { return isalnum((unsigned char)c) || c == '_'; };
    size_t pos = 0;
    while ((pos = text.find(word, pos)) != std::string::npos) {
        bool lb = pos == 0 || !isw(text[pos - 1]);
        bool rb = pos + word.size() >= text.size() || !isw(text[pos + word.size()]);
        if (lb && rb) return true;
        pos++;
    }
    return false;
}
int dx_parse_transform(const std::string& text0, ParsedTransform* out, std::string* err) {
    std::vector<std::string> buffer; std::string table_name; bool has_name = false;
    std::vector<std::string> order;
    auto append = [&](bool named) -> bool {
        std::string sql;
        for (auto& l : buffer) { if (l.empty()) continue; if (!sql.empty()) sql += " "; sql += l; }
        SqlCommand c; c.text = sql; c.name = named ? table_name : ""; c.is_query = named;
        out->commands.push_back(c);
        if (named && !table_name.empty()) {
            if (out->view_reference_count.count(table_name)) {
                *err = "dataset name '" + table_name + "' has been created, please check the query to make sure it is not created again";
                return false;
            }
            out->view_reference_count[table_name] = 0;
            for (auto& kv : out->view_reference_count) if (word_boundary_contains(sql, kv.first)) kv.second++;
        }
        return true;
    };
    for (auto& line : split_lines(strip_bom(text0))) {
        if (trim(line).empty()) continue;
        if (line == "--DataXQuery--") {               // String.matches("^--DataXQuery--"): the whole line
            if (!buffer.empty()) { if (!append(has_name)) return DX_ERR_PARSE; }
            has_name = false; table_name.clear(); buffer.clear();
            continue;
        }
        size_t k = 0; while (k < line.size() && isspace((unsigned char)line[k])) k++;
        if (line.compare(k, 2, "--") == 0) continue;  // comment line
        if (buffer.empty()) {
            // ^\s*([a-zA-Z0-9_]+)\s*=(.*)$
            size_t a = k, b = a;
            while (b < line.size() && (isalnum((unsigned char)line[b]) || line[b] == '_')) b++;
            size_t c = b; while (c < line.size() && isspace((unsigned char)line[c])) c++;
            if (b > a && c < line.size() && line[c] == '=') {
                table_name = line.substr(a, b - a); has_name = true;
                buffer.push_back(trim(line.substr(c + 1)));
            } else buffer.push_back(trim(line));
        } else buffer.push_back(trim(line));
    }
    if (has_name && !buffer.empty()) { if (!append(true)) return DX_ERR_PARSE; }
    return DX_OK;
}

// ---- stringToTimestamp (reference: datax/utility/ConcurrentDateFormat.scala:14-62), UTC session ----
static long long days_from_civil(long long y, long long m, long long d) {
    y -= m <= 2;
    const long long era = (y >= 0 ? y : y - 399) / 400;
    const long long yoe = y - era * 400;
    const long long doy = (153 * (m + (m > 2 ? -3 : 9)) + 2) / 5 + d - 1;
    const long long doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
    return era * 146097 + doe - 719468;
}
static long long lenient_epoch_s(long long y, long long mo, long long d, long long h, long long mi, long long s) {
    long long m0 = mo - 1;
    long long yy = y + (m0 >= 0 ? m0 / 12 : -((11 - m0) / 12));
    long long mm = ((m0 % 12) + 12) % 12 + 1;
    return (days_from_civil(yy, mm, 1) + (d - 1)) * 86400 + h * 3600 + mi * 60 + s;
}
static bool parse_uint(const std::string& s, size_t& i, long long* v, int min_digits = 1, int max_digits = 18) {
    size_t j = i; long long x = 0;
    while (j < s.size() && isdigit((unsigned char)s[j]) && (int)(j - i) < max_digits) { x = x * 10 + (s[j] - '0'); j++; }
    if ((int)(j - i) < min_digits) return false;
    *v = x; i = j; return true;
}
static bool timestamp_value_of(const std::string& s0, long long* us) {      // java.sql.Timestamp.valueOf
    std::string s = trim(s0);
    size_t sp = s.find(' ');
    if (sp == std::string::npos) return false;
    std::string d = s.substr(0, sp), t = s.substr(sp + 1);
    size_t d1 = d.find('-'), d2 = d1 == std::string::npos ? d1 : d.find('-', d1 + 1);
    if (d1 == std::string::npos || d2 == std::string::npos || d1 != 4 || d2 - d1 - 1 < 1 || d2 - d1 - 1 > 2 ||
        d.size() - d2 - 1 < 1 || d.size() - d2 - 1 > 2) return false;
    for (size_t i = 0; i < d.size(); i++) if (i != d1 && i != d2 && !isdigit((unsigned char)d[i])) return false;
    long long y = atoll(d.substr(0, 4).c_str()), mo = atoll(d.substr(d1 + 1, d2 - d1 - 1).c_str()), da = atoll(d.substr(d2 + 1).c_str());
    size_t c1 = t.find(':'), c2 = c1 == std::string::npos ? c1 : t.find(':', c1 + 1);
    if (c1 == std::string::npos || c2 == std::string::npos || c1 == 0 || c2 == c1 + 1 || c2 + 1 >= t.size()) return false;
    size_t p = t.find('.', c2 + 1);
    std::string hs = t.substr(0, c1), ms = t.substr(c1 + 1, c2 - c1 - 1);
    std::string ss = p == std::string::npos ? t.substr(c2 + 1) : t.substr(c2 + 1, p - c2 - 1);
    std::string fs = p == std::string::npos ? "" : t.substr(p + 1);
    auto alld = [](const std::string& x) { if (x.empty()) return false; for (char c : x) if (!isdigit((unsigned char)c)) return false; return true; };
    if (!alld(hs) || !alld(ms) || !alld(ss)) return false;
    if (p != std::string::npos && (!alld(fs) || fs.size() > 9)) return false;
    long long nanos = 0;
    if (!fs.empty()) { std::string f = fs; while (f.size() < 9) f.push_back('0'); nanos = atoll(f.c_str()); }
    *us = lenient_epoch_s(y, mo, da, atoll(hs.c_str()), atoll(ms.c_str()), atoll(ss.c_str())) * 1000000 + nanos / 1000;
    return true;
}
static bool simple_date_format(const std::string& s, const char* pattern, long long* us) {   // lenient SimpleDateFormat.parse
    long long y = 1970, mo = 1, d = 1, h = 0, mi = 0, se = 0;
    size_t i = 0;
    while (i < s.size() && isspace((unsigned char)s[i])) i++;   // DateFormat.parse skips nothing; kept strict below
    i = 0;
    for (const char* p = pattern; *p;) {
        char c = *p;
        if (c == '\'') { p++; while (*p && *p != '\'') { if (i >= s.size() || s[i] != *p) return false; i++; p++; } if (*p) p++; continue; }
        if (c == 'y' || c == 'M' || c == 'd' || c == 'H' || c == 'm' || c == 's') {
            while (*p == c) p++;
            long long v; if (!parse_uint(s, i, &v)) return false;
            if (c == 'y') y = v; else if (c == 'M') mo = v; else if (c == 'd') d = v; else if (c == 'H') h = v; else if (c == 'm') mi = v; else se = v;
            continue;
        }
        if (i >= s.size() || s[i] != c) return false;
        i++; p++;
    }
    *us = lenient_epoch_s(y, mo, d, h, mi, se) * 1000000;
    return true;
}
bool dx_string_to_timestamp_us(const std::string& s, long long* out_us) {
    if (s.empty()) return false;
    if (timestamp_value_of(s, out_us)) return true;
    if (simple_date_format(s, "yyyy-MM-dd'T'HH:mm:ss'Z'", out_us)) return true;
    if (simple_date_format(s, "MM/dd/yyyy HH:mm:ss", out_us)) return true;
    return false;
}

// =======================================================================================
// table helpers
// =======================================================================================
static std::string last_part(const std::string& n) { size_t k = n.rfind('.'); return k == std::string::npos ? n : n.substr(k + 1); }

__global__ void k_fill_i64(long long* p, long long n, long long v) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_minmax_i64(const long long* v, const void* vmask, int vwidth, int vbit, long long n, long long cut, long long* out /* [min, max, count] over the non-NULL values >= cut */) {
    long long mn = 0x7FFFFFFFFFFFFFFFll, mx = (long long)0x8000000000000000ull, cnt = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        bool valid = true;
        if (vmask) { uint32_t m = vwidth == 1 ? reinterpret_cast<const uint8_t*>(vmask)[i] : reinterpret_cast<const uint32_t*>(vmask)[i]; valid = (m >> vbit) & 1u; }
        if (valid) { long long x = v[i]; if (x >= cut) { mn = x < mn ? x : mn; mx = x > mx ? x : mx; cnt++; } }
    }
    for (int s = 16; s > 0; s >>= 1) {
        long long a = __shfl_xor_sync(0xffffffffu, mn, s), b = __shfl_xor_sync(0xffffffffu, mx, s);
        mn = a < mn ? a : mn; mx = b > mx ? b : mx; cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
    }
    // one atomic triple per CTA: thousands of warps on the same three addresses serialise in L2
    __shared__ long long s_mn[8], s_mx[8], s_cnt[8];
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) { s_mn[w] = mn; s_mx[w] = mx; s_cnt[w] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < (int)(blockDim.x >> 5); k++) { mn = s_mn[k] < mn ? s_mn[k] : mn; mx = s_mx[k] > mx ? s_mx[k] : mx; cnt += s_cnt[k]; }
        if (cnt) { atomicMin(&out[0], mn); atomicMax(&out[1], mx); atomicAdd(reinterpret_cast<unsigned long long*>(&out[2]), (unsigned long long)cnt); }
    }
}
// every output column of one group-by in one launch: the key-null word is read once per row and the validity bits of the columns
// that share a mask word leave as one reduction.  mode 0: key copy (valid = !keynull bit); 1: accumulator copy, always valid;
// 2: accumulator copy, valid iff cnt > 0; 3: average = sum / cnt
#define FIN_MAX 16
struct FinJobs {
    int n;
    int mode[FIN_MAX], keybit[FIN_MAX], width[FIN_MAX], vbit[FIN_MAX];
    const unsigned long long* a[FIN_MAX]; const unsigned long long* cnt[FIN_MAX]; void* out[FIN_MAX];
};
__global__ void __launch_bounds__(256) k_finalize_multi(const __grid_constant__ FinJobs J, const unsigned long long* __restrict__ keynull, long long n, uint32_t* __restrict__ vmask32) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long kn = keynull[i];
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < FIN_MAX; j++) {
            if (j >= J.n) break;
            bool valid = true; unsigned long long v = J.a[j] ? J.a[j][i] : 0;
            const int mode = J.mode[j];
            if (mode == 0) valid = !((kn >> J.keybit[j]) & 1ull);
            else if (mode == 2) valid = J.cnt[j][i] > 0;
            else if (mode == 3) { const unsigned long long c = J.cnt[j][i]; valid = c > 0; double s = __longlong_as_double((long long)v); v = valid ? (unsigned long long)__double_as_longlong(s / (double)c) : 0ull; }
            if (!valid) v = 0;
            if (J.width[j] == 4) reinterpret_cast<int32_t*>(J.out[j])[i] = (int32_t)v; else reinterpret_cast<unsigned long long*>(J.out[j])[i] = v;
            if (valid) bits |= 1u << J.vbit[j];
        }
        if (bits) atomicOr(&vmask32[i], bits);
    }
}

// Allocates the validity mask arrays for a freshly materialised table: one uint32 per row per
// group of 32 columns, zeroed.
struct MaskAlloc {
    dx_ctx* ctx; int64_t n; std::vector<DevBufP> bufs;
    MaskAlloc(dx_ctx* c, int64_t n_) : ctx(c), n(n_) {}
    int assign(DxColumn* col, int index) {
        size_t g = (size_t)index / 32;
        while (bufs.size() <= g) {
            DevBufP b = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 4);
            if (!b->p) return ctx->set_error(DX_ERR_CUDA, "mask allocation failed");
            DX_CUDA(ctx, cudaMemsetAsync(b->p, 0, (size_t)(n > 0 ? n : 1) * 4, ctx->stream));
            bufs.push_back(b);
        }
        col->vmask = bufs[g]; col->vwidth = 4; col->vbit = index % 32;
        return DX_OK;
    }
};

static ExprP ts_const(long long us) { auto e = std::make_shared<Expr>(); e->kind = EK::Func; e->sval = "__TS_US"; e->ival = us; return e; }

// =======================================================================================
// statement execution
// =======================================================================================
struct Exec {
    dx_flow* F; dx_ctx* ctx;
    ViewMap* views;
    long long now_us;
    std::string stmt_name;

    int fail(int code, const std::string& m) { return ctx->set_error(code, (stmt_name.empty() ? std::string() : "[" + stmt_name + "] ") + m); }
    // Lookup tables are cached under their source dictionary's address and hold that dictionary alive; one whose dictionary
    // nobody else references any more (a transient UNION / constant dictionary, a replaced state table) can never be asked
    // for again and is dropped here, so the cache does not grow with the number of micro-batches.
    void prune_luts() {
        for (auto it = F->luts.begin(); it != F->luts.end(); ) {
            if (it->second && it->second->src_dict && it->second->src_dict.use_count() == 1) it = F->luts.erase(it); else ++it;
        }
    }

    BindScope scope_for(const DxTable* t, const SelectCore& core) {
        BindScope sc; sc.table = t; sc.now_us = now_us;
        sc.qualifiers.push_back(core.from);
        if (!core.from_alias.empty()) sc.qualifiers.push_back(core.from_alias);
        return sc;
    }

    // WHERE (plus the view's time range) -> selection indices for one chunk.  sel == null && !filtered: all rows.
    int chunk_selection(const View& v, const DxTableP& chunk, const SelectCore& core, DevBufP* sel, int64_t* n_sel, bool* filtered) {
        ExprP cond = core.where;
        if (v.ranged) {
            auto col = std::make_shared<Expr>(); col->kind = EK::Col; col->parts = {v.ts_col};
            auto ge = std::make_shared<Expr>(); ge->kind = EK::Bin; ge->sval = ">="; ge->args = {col, ts_const(chunk->has_cut ? std::max(v.lo_us, chunk->ts_cut) : v.lo_us)};
            auto lt = std::make_shared<Expr>(); lt->kind = EK::Bin; lt->sval = "<"; lt->args = {col, ts_const(v.hi_us)};
            auto rng = std::make_shared<Expr>(); rng->kind = EK::Bin; rng->sval = "AND"; rng->args = {ge, lt};
            if (cond) { auto both = std::make_shared<Expr>(); both->kind = EK::Bin; both->sval = "AND"; both->args = {rng, cond}; cond = both; }
            else cond = rng;
        }
        if (v.ranged && chunk->has_ts_bounds && (chunk->ts_max < v.lo_us || chunk->ts_min >= v.hi_us)) {
            *filtered = true; *n_sel = 0; *sel = make_buf(ctx, 8); return DX_OK;      // no row of this cached batch is inside the window
        }
        if (!cond) { *filtered = false; *n_sel = chunk->n_rows; return DX_OK; }
        *filtered = true;
        if (chunk->n_rows == 0) { *n_sel = 0; *sel = make_buf(ctx, 8); return DX_OK; }
        ExprPrograms P; BindScope sc = scope_for(chunk.get(), core);
        int rc = dx_compile_expr(ctx, sc, cond, &P); if (rc) return rc;
        DevPrograms D; rc = dx_upload_programs(ctx, &P, &D); if (rc) return rc;
        DevBuf flags(ctx, (size_t)chunk->n_rows * 4);
        rc = dx_eval_flags(ctx, D, chunk->n_rows, nullptr, flags.as<uint32_t>(), 1); if (rc) return rc;
        return dx_select_rows(ctx, flags.as<uint32_t>(), 1, chunk->n_rows, sel, n_sel);
    }

    static bool is_agg_func(const ExprP& e) {
        if (e->kind != EK::Func) return false;
        const std::string& f = e->sval;
        return f == "COUNT" || f == "SUM" || f == "MIN" || f == "MAX" || f == "AVG";
    }
    static bool contains_agg(const ExprP& e) {
        if (!e) return false;
        if (is_agg_func(e)) return true;
        for (auto& a : e->args) if (contains_agg(a)) return true;
        return false;
    }
    static bool has_column_ref(const ExprP& e) {
        if (!e) return false;
        if (e->kind == EK::Col || e->kind == EK::Star) return true;
        for (auto& a : e->args) if (has_column_ref(a)) return true;
        return false;
    }
    static bool is_literal(const ExprP& e) { return e->kind == EK::Int || e->kind == EK::Dec || e->kind == EK::Str || e->kind == EK::Null || e->kind == EK::Bool; }

    static std::string item_name(const SelectItem& it) {
        if (!it.alias.empty()) return it.alias;
        if (it.expr->kind == EK::Col) return it.expr->parts.back();
        if (it.expr->kind == EK::Str) return it.expr->sval;
        return to_string(it.expr);
    }
    static DxColumn literal_column(const ExprP& e, const std::string& name) {
        DxColumn c; c.name = name; c.is_const = true;
        switch (e->kind) {
        case EK::Int: c.type = DX_T_I64; c.const_i = e->ival; break;
        case EK::Dec: c.type = DX_T_F64; c.const_f = e->dval; break;
        case EK::Bool: c.type = DX_T_BOOL; c.const_i = e->ival; break;
        case EK::Str: c.type = DX_T_STR; c.const_s = e->sval; c.const_i = -1; break;
        default: c.type = DX_T_NULL; c.all_null = true; break;
        }
        return c;
    }

    // filterNull(Array(IF(cond, MAP('k','v',...), NULL), ...)) and IF(cond, MAP(...), NULL):
    // the rule-array shape emitted by CodegenRules (Engine.cs:232-266, 380-397; Rule.cs:272-281).
    static bool match_rule_if(const ExprP& e, ExprP* cond, std::vector<std::pair<std::string, std::string>>* map) {
        if (e->kind != EK::Func || e->sval != "IF" || e->args.size() != 3) return false;
        if (e->args[2]->kind != EK::Null) return false;
        const ExprP& m = e->args[1];
        if (m->kind != EK::Func || m->sval != "MAP" || m->args.size() % 2) return false;
        for (size_t i = 0; i < m->args.size(); i += 2) {
            if (m->args[i]->kind != EK::Str) return false;
            const ExprP& v = m->args[i + 1];
            std::string sv;
            if (v->kind == EK::Str) sv = v->sval; else if (v->kind == EK::Int) sv = std::to_string(v->ival); else if (v->kind == EK::Dec) sv = v->sval; else return false;
            map->emplace_back(m->args[i]->sval, sv);
        }
        *cond = e->args[0];
        return true;
    }
    static bool match_rules(const ExprP& e, std::vector<ExprP>* conds, std::shared_ptr<RuleTable>* rt) {
        auto table = std::make_shared<RuleTable>();
        ExprP c; std::vector<std::pair<std::string, std::string>> m;
        if (match_rule_if(e, &c, &m)) { conds->push_back(c); table->rules.push_back(m); table->single = true; *rt = table; return true; }
        if (e->kind != EK::Func || upper(e->sval) != "FILTERNULL" || e->args.size() != 1) return false;
        const ExprP& arr = e->args[0];
        if (arr->kind != EK::Func || arr->sval != "ARRAY") return false;
        for (auto& a : arr->args) {
            m.clear();
            if (!match_rule_if(a, &c, &m)) return false;
            conds->push_back(c); table->rules.push_back(m);
        }
        *rt = table;
        return !conds->empty();
    }

    // UNION ALL whose branches only forward columns (the accumulator pattern: SELECT cols FROM new UNION ALL SELECT cols
    // FROM state WHERE retention, StateTableHandler / UserCodeIoTSample.txt): every branch's selected rows are gathered
    // ONCE, straight into their slice of the result, instead of materialising each branch and concatenating the copies.
    // *done = false: some branch needs the general path.
    int exec_union_all_direct(const Statement& st, DxTableP* out, bool* done) {
        *done = false;
        struct Branch { const DxTable* src; std::vector<int> cols; DevBufP sel; int64_t n = 0; bool filtered = false; };
        std::vector<Branch> br(st.selects.size());
        std::vector<std::string> names;
        for (bool all : st.union_all) if (!all) return DX_OK;
        for (size_t k = 0; k < st.selects.size(); k++) {
            const SelectCore& core = st.selects[k];
            if (!core.joins.empty() || core.distinct || !core.group_by.empty()) return DX_OK;
            auto vit = views->find(core.from);
            if (vit == views->end() || vit->second.chunks.size() != 1) return DX_OK;
            const View& v = vit->second;
            const DxTable* t = v.chunks[0].get();
            BindScope sc = scope_for(t, core); sc.qualifiers.push_back("Raw");
            for (auto& it : core.items) {
                if (it.expr->kind == EK::Star && it.expr->parts.empty()) {
                    for (size_t c = 0; c < t->cols.size(); c++) { br[k].cols.push_back((int)c); if (k == 0) names.push_back(t->cols[c].name); }
                } else if (it.expr->kind == EK::Col) {
                    std::string err; int ci = dx_resolve_column(sc, it.expr->parts, &err);
                    if (ci < 0) return DX_OK;
                    br[k].cols.push_back(ci); if (k == 0) names.push_back(item_name(it));
                } else return DX_OK;
            }
            if (br[k].cols.size() != br[0].cols.size()) return DX_OK;
            for (size_t c = 0; c < br[k].cols.size(); c++) {
                const DxColumn& a0 = br[0].src ? br[0].src->cols[br[0].cols[c]] : t->cols[br[k].cols[c]];
                const DxColumn& ck = t->cols[br[k].cols[c]];
                if (!ck.data || ck.all_null || ck.type == DX_T_FLAGS || ck.type == DX_T_NULL || !ck.nest_path.empty()) return DX_OK;
                if (k > 0 && (a0.type != ck.type || (ck.type == DX_T_STR && a0.dict != ck.dict))) return DX_OK;
            }
            br[k].src = t;
            int rc = chunk_selection(v, v.chunks[0], core, &br[k].sel, &br[k].n, &br[k].filtered); if (rc) return rc;
        }
        int64_t total = 0; for (auto& b : br) total += b.n;
        auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = total;
        MaskAlloc masks(ctx, total);
        const size_t nc = br[0].cols.size();
        for (size_t c = 0; c < nc; c++) {
            const DxColumn& proto = br[0].src->cols[br[0].cols[c]];
            DxColumn oc; oc.name = names[c]; oc.type = proto.type; oc.dict = proto.dict;
            const int w = dx_type_width(oc.type);
            oc.data = make_buf(ctx, (size_t)(total > 0 ? total : 1) * w);
            if (!oc.data->p) return fail(DX_ERR_CUDA, "UNION ALL allocation failed");
            int rc = masks.assign(&oc, (int)c); if (rc) return rc;
            int64_t off = 0;
            for (auto& b : br) {
                if (b.n == 0) continue;
                rc = dx_gather_column(ctx, b.src->cols[b.cols[c]], b.n, b.filtered ? b.sel->as<int64_t>() : nullptr,
                                      oc.data->as<uint8_t>() + (size_t)off * w, oc.vmask->as<uint32_t>() + off, oc.vbit);
                if (rc) return rc;
                off += b.n;
            }
            t->cols.push_back(oc);
        }
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        *out = t; *done = true;
        return DX_OK;
    }

    int exec_statement(const Statement& st, DxTableP* out) {
        if (st.selects.size() > 1) {
            bool done = false;
            int rc = exec_union_all_direct(st, out, &done); if (rc) return rc;
            if (done) return DX_OK;
        }
        std::vector<DxTableP> parts;
        for (auto& core : st.selects) { DxTableP t; int rc = exec_core(core, &t); if (rc) return rc; parts.push_back(t); }
        if (parts.size() == 1) { *out = parts[0]; return DX_OK; }
        DxTableP u; int rc = concat(parts, &u); if (rc) return rc;
        bool any_distinct = false; for (bool a : st.union_all) if (!a) any_distinct = true;
        if (!any_distinct) { *out = u; return DX_OK; }
        // UNION = UNION ALL + DISTINCT over all columns
        View v; v.chunks.push_back(u);
        (*views)["__union"] = v;
        SelectCore d; d.distinct = true; d.from = "__union";
        SelectItem star; star.expr = std::make_shared<Expr>(); star.expr->kind = EK::Star; d.items.push_back(star);
        rc = exec_core(d, out);
        views->erase("__union");
        return rc;
    }

    // positional concatenation (UNION ALL): same arity, compatible physical types
    int concat(const std::vector<DxTableP>& parts, DxTableP* out) {
        size_t nc = parts[0]->cols.size(); int64_t total = 0;
        for (auto& p : parts) { if (p->cols.size() != nc) return fail(DX_ERR_UNSUPPORTED, "UNION branches have different column counts"); total += p->n_rows; }
        auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = total;
        MaskAlloc masks(ctx, total);
        for (size_t c = 0; c < nc; c++) {
            DxColumn oc; oc.name = parts[0]->cols[c].name; oc.nest_path = parts[0]->cols[c].nest_path; oc.nest_kinds = parts[0]->cols[c].nest_kinds;
            int type = DX_T_NULL; DxDictP dict; bool all_const_same = true;
            for (auto& p : parts) {
                const DxColumn& pc = p->cols[c];
                if (pc.type == DX_T_NULL || pc.all_null) continue;
                if (type == DX_T_NULL) type = pc.type;
                else if (type != pc.type) {
                    bool il = (type == DX_T_I64 || type == DX_T_BOOL) && (pc.type == DX_T_I64 || pc.type == DX_T_BOOL);
                    // Spark widens timestamp + string to string (TypeCoercion: string promotion)
                    const bool ts_str = (type == DX_T_STR && pc.type == DX_T_TS) || (type == DX_T_TS && pc.type == DX_T_STR);
                    if (ts_str) type = DX_T_STR;
                    else if (!il) return fail(DX_ERR_UNSUPPORTED, "UNION column '" + oc.name + "' mixes types");
                }
                if (pc.type == DX_T_STR && pc.data && p->n_rows > 0) {
                    if (!dict) dict = pc.dict;      // codes of the other sides are re-encoded into the first side's dictionary
                }
                if (pc.type == DX_T_FLAGS) return fail(DX_ERR_UNSUPPORTED, "UNION over a rule-flag column");
            }
            (void)all_const_same;
            oc.type = type;
            if (type == DX_T_NULL) { oc.all_null = true; t->cols.push_back(oc); continue; }
            if (type == DX_T_STR && !dict) {
                // only constant strings on every side (e.g. the rule constants of two aggregate rules): one shared dictionary
                dict = std::make_shared<DxDict>();
                int rcd = dict->init(ctx, 1024, 1 << 16); if (rcd) return rcd;
            }
            oc.dict = dict;
            int w = dx_type_width(type);
            oc.data = make_buf(ctx, (size_t)(total > 0 ? total : 1) * w);
            int rc = masks.assign(&oc, (int)c); if (rc) return rc;
            int64_t off = 0;
            for (auto& p : parts) {
                const DxColumn& pc = p->cols[c];
                if (p->n_rows == 0) continue;
                uint8_t* dst = oc.data->as<uint8_t>() + (size_t)off * w;
                uint32_t* vm = oc.vmask->as<uint32_t>() + off;
                if (pc.type == DX_T_NULL || pc.all_null) { DX_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t)p->n_rows * w, ctx->stream)); }
                else if (type == DX_T_STR && pc.type == DX_T_TS) {
                    // this side's timestamps as text, entered into the result's dictionary
                    DxColumn src = pc;
                    if (!src.data) { rc = materialize_const(&src, p->n_rows, nullptr); if (rc) return rc; }
                    StrBuildArgs sa; memset(&sa, 0, sizeof sa);
                    sa.n_parts = 1; sa.parts[0].kind = SP_TS;
                    sa.parts[0].col.data = src.data->p; sa.parts[0].col.vmask = src.vmask ? src.vmask->p : nullptr;
                    sa.parts[0].col.vwidth = src.vwidth; sa.parts[0].col.vbit = src.vbit; sa.parts[0].col.width = 8;
                    sa.n_rows = p->n_rows; sa.sel = nullptr;
                    rc = dx_strbuild(ctx, sa, std::string(), dict, reinterpret_cast<int32_t*>(dst), vm, oc.vbit); if (rc) return rc;
                }
                else if (!pc.data) {
                    DxColumn tmp = pc;
                    rc = materialize_const(&tmp, p->n_rows, dict); if (rc) return rc;
                    rc = dx_gather_column(ctx, tmp, p->n_rows, nullptr, dst, vm, oc.vbit); if (rc) return rc;
                } else if (pc.type == DX_T_STR && pc.dict && pc.dict != dict) {
                    // another dictionary (a restored state table, a CONCAT result ...): translate code -> code through a
                    // per-(source, target) lookup table that only grows with the source dictionary
                    char kb[64]; snprintf(kb, sizeof kb, "recode@%p/%p", (void*)pc.dict.get(), (void*)dict.get());
                    std::shared_ptr<LutCache>& lut = F->luts[kb];
                    if (!lut) { prune_luts(); lut = std::make_shared<LutCache>(); lut->src_dict = pc.dict; lut->out_dict = dict; }
                    DxDictP target = dict;
                    rc = dx_lut_refresh(ctx, lut.get(), pc.dict, [&](const std::string& sv, long long* v) {
                        int32_t code = -1; if (target->insert_host(sv, &code)) return false; *v = code; return true; });
                    if (rc) return rc;
                    rc = dx_lut_apply(ctx, pc, *lut, p->n_rows, nullptr, dst, 4, vm, oc.vbit); if (rc) return rc;
                } else { rc = dx_gather_column(ctx, pc, p->n_rows, nullptr, dst, vm, oc.vbit); if (rc) return rc; }
                off += p->n_rows;
            }
            t->cols.push_back(oc);
        }
        *out = t; return DX_OK;
    }

    // turns a constant column into a materialised one (needed when it meets real columns in a UNION / as a group key)
    int materialize_const(DxColumn* c, int64_t n, DxDictP dict) {
        long long v = c->const_i;
        if (c->type == DX_T_F64) memcpy(&v, &c->const_f, 8);
        if (c->type == DX_T_STR) {
            if (!dict) { dict = std::make_shared<DxDict>(); int rc = dict->init(ctx, 1024, 1 << 16); if (rc) return rc; }
            int32_t code; int rc = dict->insert_host(c->const_s, &code); if (rc) return rc;
            v = code; c->dict = dict;
        }
        // physical width 8 for the fill, then narrowed for dictionary codes
        DevBufP b = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 8);
        if (n > 0) { k_fill_i64<<<(unsigned)std::min<long long>((n + 255) / 256, 4096), 256, 0, ctx->stream>>>(b->as<long long>(), n, v); ctx->launches["k_fill"]++; }
        if (c->type == DX_T_STR && n > 0) {
            DevBufP b4 = make_buf(ctx, (size_t)n * 4);
            DxColumn wide; wide.type = DX_T_I64; wide.data = b;
            // narrow by gathering through the 8-byte view
            std::vector<int32_t> host((size_t)n, (int32_t)v);
            DX_CUDA(ctx, cudaMemcpyAsync(b4->p, host.data(), (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
            DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            b = b4;
        }
        c->data = b; c->is_const = false; c->vmask = nullptr;
        return DX_OK;
    }

    // MAP('k', v, ...) / STRUCT(v AS n, ...) / NAMED_STRUCT('n', v, ...) select items — the alert shaping
    // CodegenRules emits (pivots / aggs / result / RuleObject) — become one column per leaf, named
    // <alias>.<key>..., with the key path and container kinds kept on the column for the sink.
    static bool is_nested_ctor(const ExprP& e) {
        return e->kind == EK::Func && (e->sval == "MAP" || e->sval == "STRUCT" || e->sval == "NAMED_STRUCT");
    }
    struct NestLeaf { std::vector<std::string> path; std::string kinds; };
    int flatten_nested(const ExprP& e, const std::string& prefix, std::vector<std::string>& path, std::string& kinds,
                       std::vector<SelectItem>* items, std::map<std::string, NestLeaf>* nest) {
        const bool is_struct = e->sval == "STRUCT";
        if (!is_struct && e->args.size() % 2) return fail(DX_ERR_ARG, e->sval + " needs key/value pairs");
        for (size_t i = 0; i < e->args.size(); i += is_struct ? 1 : 2) {
            std::string key; ExprP val;
            if (is_struct) {
                val = e->args[i];
                key = !val->alias.empty() ? val->alias : (val->kind == EK::Col ? val->parts.back() : "col" + std::to_string(i + 1));
            } else {
                if (e->args[i]->kind != EK::Str) return fail(DX_ERR_UNSUPPORTED, e->sval + " keys must be string literals");
                key = e->args[i]->sval; val = e->args[i + 1];
            }
            path.push_back(key); kinds.push_back(is_struct ? 's' : 'm');
            const std::string name = prefix + "." + key;
            if (is_nested_ctor(val)) { int rc = flatten_nested(val, name, path, kinds, items, nest); if (rc) return rc; }
            else {
                SelectItem it; it.expr = val; it.alias = name; items->push_back(it);
                NestLeaf nl; nl.path = path; nl.kinds = kinds; (*nest)[name] = nl;
            }
            path.pop_back(); kinds.pop_back();
        }
        return DX_OK;
    }

    // HAVING cond  ==  the aggregation with every aggregate (and bare column) of cond added as a hidden item, then a WHERE over
    // its result that keeps the visible items only.
    ExprP having_rewrite(const ExprP& e, std::vector<SelectItem>* inner, const std::vector<SelectItem>& visible) {
        if (!e) return e;
        if (is_agg_func(e) || e->kind == EK::Col) {
            const std::string key = to_string(e);
            for (auto& it : *inner) if (to_string(it.expr) == key) {
                auto c = std::make_shared<Expr>(); c->kind = EK::Col; c->parts = {item_name(it)}; return c;
            }
            if (e->kind == EK::Col && e->parts.size() == 1)          // a select alias
                for (auto& it : visible) if (item_name(it) == e->parts[0]) return e;
            SelectItem h; h.expr = e; h.alias = "__having" + std::to_string(inner->size());
            inner->push_back(h);
            auto c = std::make_shared<Expr>(); c->kind = EK::Col; c->parts = {h.alias}; return c;
        }
        auto cp = std::make_shared<Expr>(*e);
        for (auto& a2 : cp->args) a2 = having_rewrite(a2, inner, visible);
        return cp;
    }

    int exec_core(const SelectCore& core, DxTableP* out) {
        if (core.having) {
            for (auto& it : core.items) if (it.expr->kind == EK::Star) return fail(DX_ERR_UNSUPPORTED, "HAVING with SELECT *");
            SelectCore inner = core; inner.having = nullptr;
            ExprP cond = having_rewrite(core.having, &inner.items, core.items);
            DxTableP agg; int rc = exec_core(inner, &agg); if (rc) return rc;
            View hv; hv.chunks.push_back(agg);
            (*views)["__having"] = hv;
            SelectCore outer; outer.from = "__having"; outer.where = cond;
            for (auto& it : core.items) {
                SelectItem oi; oi.expr = std::make_shared<Expr>(); oi.expr->kind = EK::Col; oi.expr->parts = {item_name(it)}; oi.alias = item_name(it);
                outer.items.push_back(oi);
            }
            rc = exec_core(outer, out);
            views->erase("__having");
            return rc;
        }
        bool nested = false;
        for (auto& it : core.items) if (is_nested_ctor(it.expr)) nested = true;
        if (nested) {
            SelectCore flat = core; flat.items.clear();
            std::map<std::string, NestLeaf> nest;
            for (auto& it : core.items) {
                if (!is_nested_ctor(it.expr)) { flat.items.push_back(it); continue; }
                if (it.alias.empty()) return fail(DX_ERR_UNSUPPORTED, "a MAP/STRUCT select item needs an alias");
                std::vector<std::string> path; std::string kinds;
                int rc = flatten_nested(it.expr, it.alias, path, kinds, &flat.items, &nest); if (rc) return rc;
            }
            int rc = exec_core(flat, out); if (rc) return rc;
            for (auto& c : (*out)->cols) {
                auto f = nest.find(c.name);
                if (f != nest.end()) { c.nest_path = f->second.path; c.nest_kinds = f->second.kinds; }
            }
            return DX_OK;
        }
        auto vit = views->find(core.from);
        if (vit == views->end()) return fail(DX_ERR_ARG, "table or view '" + core.from + "' not found");
        if (!core.joins.empty()) return exec_join(core, out);
        const View& v = vit->second;
        if (v.chunks.empty()) return fail(DX_ERR_STATE, "view '" + core.from + "' has no data");
        bool agg = core.distinct || !core.group_by.empty();
        for (auto& it : core.items) if (contains_agg(it.expr)) agg = true;
        return agg ? exec_agg(core, v, out) : exec_plain(core, v, out);
    }

    // ---------------- [INNER] JOIN ... ON a = b [AND c = d] ----------------
    int materialize_view(const std::string& name, DxTableP* out) {
        auto vit = views->find(name);
        if (vit == views->end()) return fail(DX_ERR_ARG, "table or view '" + name + "' not found");
        const View& v = vit->second;
        if (v.chunks.size() == 1 && !v.ranged) { *out = v.chunks[0]; return DX_OK; }
        SelectCore all; all.from = name;
        SelectItem star; star.expr = std::make_shared<Expr>(); star.expr->kind = EK::Star; all.items.push_back(star);
        return exec_plain(all, v, out);
    }
    // join key as an int64 column: integers as they are, strings parsed to long when the other side is numeric
    // (Spark casts the string side; unparsable -> NULL, which never matches), string-vs-string through the left dictionary
    int join_key_column(const DxTable* t, int ci, bool other_is_string, const DxColumn* left_str, DxColumn* out) {
        const DxColumn& c = t->cols[ci];
        if (!c.data || c.all_null) return fail(DX_ERR_UNSUPPORTED, "join key '" + c.name + "' is a constant column");
        if (c.type == DX_T_I64 || c.type == DX_T_TS || c.type == DX_T_BOOL) { *out = c; return DX_OK; }
        if (c.type != DX_T_STR) return fail(DX_ERR_UNSUPPORTED, "join on a double or flag column");
        char kb[96]; snprintf(kb, sizeof kb, "@%p/%p/%d", (void*)c.dict.get(), left_str ? (void*)left_str->dict.get() : nullptr, (int)other_is_string);
        std::shared_ptr<LutCache>& lut = F->luts[std::string("joinkey") + kb];
        if (!lut) { prune_luts(); lut = std::make_shared<LutCache>(); lut->src_dict = c.dict; }    // the key is the dictionary's address: keep it alive
        int rc;
        if (!other_is_string) {
            rc = dx_lut_refresh(ctx, lut.get(), c.dict, [](const std::string& s, long long* v) {
                if (s.empty()) return false;
                size_t a = 0, b = s.size();
                while (a < b && isspace((unsigned char)s[a])) a++;
                while (b > a && isspace((unsigned char)s[b - 1])) b--;
                if (a == b) return false;
                size_t i = a; bool neg = false;
                if (s[i] == '-' || s[i] == '+') { neg = s[i] == '-'; i++; }
                if (i == b) return false;
                long long x = 0;
                for (; i < b; i++) { if (!isdigit((unsigned char)s[i])) return false; x = x * 10 + (s[i] - '0'); }
                *v = neg ? -x : x; return true;
            });
        } else if (left_str) {
            DxDictP ld = left_str->dict;
            rc = ld->sync_to_host(); if (rc) return rc;
            lut->built = 0;   // the left dictionary may have grown: rebuild (reference tables are small)
            rc = dx_lut_refresh(ctx, lut.get(), c.dict, [&](const std::string& s, long long* v) { int k = ld->lookup_host(s); if (k < 0) return false; *v = k; return true; });
        } else { *out = c; return DX_OK; }     // the left side of a string-vs-string key: its own codes
        if (rc) return rc;
        out->name = c.name; out->type = DX_T_I64; out->is_const = false; out->all_null = false;
        out->data = make_buf(ctx, (size_t)(t->n_rows > 0 ? t->n_rows : 1) * 8);
        out->vmask = make_buf(ctx, (size_t)(t->n_rows > 0 ? t->n_rows : 1) * 4); out->vwidth = 4; out->vbit = 0;
        DX_CUDA(ctx, cudaMemsetAsync(out->vmask->p, 0, (size_t)(t->n_rows > 0 ? t->n_rows : 1) * 4, ctx->stream));
        return dx_lut_apply(ctx, c, *lut, t->n_rows, nullptr, out->data->p, 8, out->vmask->as<uint32_t>(), 0);
    }
    static void flatten_and(const ExprP& e, std::vector<ExprP>* out) {
        if (e->kind == EK::Bin && e->sval == "AND") { flatten_and(e->args[0], out); flatten_and(e->args[1], out); } else out->push_back(e);
    }
    int exec_join(const SelectCore& core, DxTableP* out) {
        DxTableP left; int rc = materialize_view(core.from, &left); if (rc) return rc;
        std::string left_name = core.from_alias.empty() ? core.from : core.from_alias;
        std::vector<std::string> owners;
        if (left->col_owner.empty()) owners.assign(left->cols.size(), left_name); else owners = left->col_owner;
        DxTableP acc = left;
        for (auto& j : core.joins) {
            DxTableP right; rc = materialize_view(j.table, &right); if (rc) return rc;
            const std::string right_name = j.alias.empty() ? j.table : j.alias;
            auto accT = std::make_shared<DxTable>(*acc); accT->col_owner = owners;      // left scope with owners for qualified names
            BindScope ls; ls.table = accT.get(); ls.now_us = now_us;
            BindScope rs; rs.table = right.get(); rs.now_us = now_us; rs.qualifiers = {j.table, right_name};
            std::vector<ExprP> conj; flatten_and(j.on, &conj);
            if (conj.empty() || conj.size() > DX_JOIN_MAX_KEYS) return fail(DX_ERR_UNSUPPORTED, "JOIN ... ON needs 1 to 4 equalities");
            JoinSide L, R; memset(&L, 0, sizeof L); memset(&R, 0, sizeof R);
            std::vector<DxColumn> keep;
            for (auto& c : conj) {
                if (c->kind != EK::Bin || c->sval != "=" || c->args[0]->kind != EK::Col || c->args[1]->kind != EK::Col)
                    return fail(DX_ERR_UNSUPPORTED, "JOIN condition '" + to_string(c) + "' is not column = column");
                std::string e1, e2;
                int la = dx_resolve_column(ls, c->args[0]->parts, &e1), rb = dx_resolve_column(rs, c->args[1]->parts, &e2);
                if (la < 0 || rb < 0) {   // written the other way round
                    la = dx_resolve_column(ls, c->args[1]->parts, &e1); rb = dx_resolve_column(rs, c->args[0]->parts, &e2);
                    if (la < 0 || rb < 0) return fail(DX_ERR_ARG, "JOIN condition '" + to_string(c) + "': " + (la < 0 ? e1 : e2));
                }
                const bool lstr = accT->cols[la].type == DX_T_STR, rstr = right->cols[rb].type == DX_T_STR;
                DxColumn lk, rk;
                rc = join_key_column(accT.get(), la, rstr, nullptr, &lk); if (rc) return rc;
                rc = join_key_column(right.get(), rb, lstr, lstr ? &accT->cols[la] : nullptr, &rk); if (rc) return rc;
                auto desc = [](const DxColumn& k) { ProgCol pc; pc.data = k.data->p; pc.vmask = k.vmask ? k.vmask->p : nullptr; pc.vwidth = k.vwidth; pc.vbit = k.vbit; pc.width = dx_type_width(k.type); pc.pad = 0; return pc; };
                L.keys[L.n_keys++] = desc(lk); R.keys[R.n_keys++] = desc(rk);
                keep.push_back(lk); keep.push_back(rk);
            }
            DevBufP li, ri; int64_t np = 0;
            rc = dx_join_pairs(ctx, L, acc->n_rows, nullptr, R, right->n_rows, &li, &ri, &np, j.left_outer); if (rc) return rc;
            auto jt = std::make_shared<DxTable>(); jt->ctx = ctx; jt->n_rows = np;
            MaskAlloc masks(ctx, np);
            std::vector<std::string> new_owners;
            int idx = 0;
            auto add_side = [&](const DxTable* t, const std::vector<std::string>& own, const DevBufP& sel, bool outer_side) -> int {
                for (size_t c = 0; c < t->cols.size(); c++) {
                    DxColumn oc = t->cols[c];
                    if (oc.type == DX_T_FLAGS) return fail(DX_ERR_UNSUPPORTED, "JOIN over a rule-flag column");
                    if (outer_side && !oc.data && !oc.all_null && oc.type != DX_T_NULL) {     // a constant would survive on unmatched rows
                        int rc3 = materialize_const(&oc, t->n_rows, nullptr); if (rc3) return rc3;
                    }
                    if (oc.data && !oc.all_null) {
                        DevBufP nb = make_buf(ctx, (size_t)(np > 0 ? np : 1) * dx_type_width(oc.type));
                        DxColumn dst = oc; dst.data = nb;
                        int rc2 = masks.assign(&dst, idx); if (rc2) return rc2;
                        if (np > 0) { rc2 = dx_gather_column(ctx, oc, np, sel->as<int64_t>(), nb->p, dst.vmask->as<uint32_t>(), dst.vbit); if (rc2) return rc2; }
                        oc = dst;
                    }
                    jt->cols.push_back(oc); new_owners.push_back(own[c]); idx++;
                }
                return DX_OK;
            };
            rc = add_side(acc.get(), owners, li, false); if (rc) return rc;
            std::vector<std::string> rown(right->cols.size(), right_name);
            rc = add_side(right.get(), rown, ri, j.left_outer); if (rc) return rc;
            DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            acc = jt; owners = new_owners;
        }
        acc->col_owner = owners;
        // the rest of the statement runs over the joined rows
        View jv; jv.chunks.push_back(acc);
        (*views)["__join"] = jv;
        SelectCore rest = core; rest.joins.clear(); rest.from = "__join"; rest.from_alias.clear();
        rc = exec_core(rest, out);
        views->erase("__join");
        if (*out) (*out)->col_owner.clear();
        return rc;
    }

    // ---------------- SELECT without aggregation: filter + project ----------------
    int exec_plain(const SelectCore& core, const View& v, DxTableP* out) {
        if (v.chunks.size() > 1) {
            // rule arrays are evaluated over one contiguous table: gather a multi-batch window view first
            bool has_rules = false;
            for (auto& it : core.items) { std::vector<ExprP> cds; std::shared_ptr<RuleTable> rt; if (match_rules(it.expr, &cds, &rt)) has_rules = true; }
            if (has_rules) {
                SelectCore all; all.from = core.from;
                SelectItem star; star.expr = std::make_shared<Expr>(); star.expr->kind = EK::Star; all.items.push_back(star);
                DxTableP m; int rc = exec_plain(all, v, &m); if (rc) return rc;
                View one; one.chunks.push_back(m);
                return exec_plain(core, one, out);
            }
        }
        struct ChunkSel { DevBufP sel; int64_t n = 0; bool filtered = false; };
        std::vector<ChunkSel> cs(v.chunks.size());
        int64_t total = 0;
        for (size_t k = 0; k < v.chunks.size(); k++) {
            int rc = chunk_selection(v, v.chunks[k], core, &cs[k].sel, &cs[k].n, &cs[k].filtered); if (rc) return rc;
            total += cs[k].n;
        }
        const bool zero_copy = v.chunks.size() == 1 && !cs[0].filtered;
        auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = total;
        MaskAlloc masks(ctx, total);
        const DxTable* proto = v.chunks[0].get();
        int out_index = 0;
        struct PendingGather { GatherJobs jobs; uint32_t* mask = nullptr; std::vector<DevBufP> keep; };
        std::vector<PendingGather> pending(v.chunks.size());
        auto flush_gather = [&](size_t k) -> int {
            PendingGather& pg = pending[k];
            if (!pg.jobs.n) return DX_OK;
            int rc = dx_gather_columns(ctx, pg.jobs, cs[k].n, cs[k].filtered ? cs[k].sel->as<int64_t>() : nullptr, pg.mask);
            pg.jobs.n = 0; pg.mask = nullptr; pg.keep.clear();
            return rc;
        };
        for (auto& it : core.items) {
            const ExprP& e = it.expr;
            std::vector<int> src_cols;   // columns forwarded as they are
            std::map<int, std::string> renames;      // leaves of a selected metadata map / struct: their names under the item
            if (e->kind == EK::Star) {
                if (e->parts.size() == 1 && e->parts[0] == "Raw" && stmt_name == "projection") {
                    // Raw.* in a projection: the decoded leaves, not the transport metadata columns that sit beside Raw
                    for (size_t c = 0; c < proto->cols.size(); c++) if (!dx_is_metadata_column(proto->cols[c].name)) src_cols.push_back((int)c);
                } else if (e->parts.empty() || e->parts[0] == core.from || e->parts[0] == core.from_alias || (e->parts.size() == 1 && e->parts[0] == "Raw"))
                    for (size_t c = 0; c < proto->cols.size(); c++) src_cols.push_back((int)c);
                else {
                    std::string prefix; for (auto& p : e->parts) prefix += p + ".";
                    if (prefix.compare(0, 4, "Raw.") == 0) prefix = prefix.substr(4);
                    for (size_t c = 0; c < proto->cols.size(); c++) if (proto->cols[c].name.compare(0, prefix.size(), prefix) == 0) src_cols.push_back((int)c);
                    if (src_cols.empty()) return fail(DX_ERR_ARG, "cannot expand '" + to_string(e) + "'");
                }
            } else if (e->kind == EK::Col) {
                BindScope sc = scope_for(proto, core); sc.qualifiers.push_back("Raw");
                std::string err; int ci = dx_resolve_column(sc, e->parts, &err);
                if (ci < 0) {
                    // a whole struct / map of transport metadata (`Properties AS _Properties`, `__DataX_FileInfo`): every leaf below
                    // it, renamed under the item's name; an empty map (or metadata this batch carries none of) is a NULL column
                    const std::string& n0 = e->parts[0];
                    std::string prefix; for (auto& pp : e->parts) prefix += pp + ".";
                    bool nested_parent = false;       // e.g. `_Properties` in a later step: the leaves `_Properties.<key>` of an earlier one
                    for (auto& pc : proto->cols) if (!pc.nest_kinds.empty() && !pc.data && pc.name.compare(0, prefix.size(), prefix) == 0) nested_parent = true;
                    if (nested_parent || n0 == "Properties" || n0 == "SystemProperties" || n0 == "__DataX_FileInfo" || n0 == "DataXProperties") {
                        // every leaf below it goes the way of a forwarded column (constants as they are, per-event values gathered
                        // like any other column), renamed under the item's name and keeping its place in the map / struct
                        for (size_t c = 0; c < proto->cols.size(); c++) {
                            if (proto->cols[c].name.compare(0, prefix.size(), prefix) != 0) continue;
                            src_cols.push_back((int)c);
                            renames[(int)c] = item_name(it) + "." + proto->cols[c].name.substr(prefix.size());
                        }
                        if (src_cols.empty()) { DxColumn c; c.name = item_name(it); c.type = DX_T_NULL; c.all_null = true; t->cols.push_back(c); out_index++; continue; }
                    } else return fail(DX_ERR_ARG, err);
                } else src_cols.push_back(ci);
            }
            if (!src_cols.empty()) {
                for (int ci : src_cols) {
                    DxColumn oc = proto->cols[ci];
                    auto rn = renames.find(ci);
                    if (rn != renames.end()) oc.name = rn->second;
                    else if (e->kind == EK::Col) { oc.name = item_name(it); oc.nest_path.clear(); oc.nest_kinds.clear(); }
                    // A per-batch constant (DataXProperties.BatchTime, a Properties value of the batch) is a different constant in
                    // every cached batch of a window view: the rows keep the value of the batch they arrived in, so the column is
                    // laid out chunk by chunk.
                    bool mixed = false; DxDictP mixed_dict;
                    if (!oc.data && v.chunks.size() > 1) {
                        for (auto& chp : v.chunks) {
                            const DxColumn& sc = chp->cols[ci];
                            if (sc.data || sc.all_null != oc.all_null || sc.type != oc.type || sc.const_i != oc.const_i || sc.const_f != oc.const_f || sc.const_s != oc.const_s) mixed = true;
                        }
                        if (mixed) {
                            int ty = DX_T_NULL;
                            for (auto& chp : v.chunks) { const DxColumn& sc = chp->cols[ci]; if (ty == DX_T_NULL && sc.type != DX_T_NULL && !sc.all_null) ty = sc.type; }
                            if (ty == DX_T_NULL) mixed = false;
                            else {
                                oc.type = ty; oc.all_null = false; oc.is_const = false;
                                if (ty == DX_T_STR) { mixed_dict = std::make_shared<DxDict>(); int rc = mixed_dict->init(ctx, 1024, 1 << 16); if (rc) return rc; oc.dict = mixed_dict; }
                            }
                        }
                    }
                    if (mixed || !(oc.all_null || !oc.data || zero_copy)) {
                        const int words = oc.type == DX_T_FLAGS ? oc.flag_words : 1;
                        const int w = oc.type == DX_T_FLAGS ? 4 : dx_type_width(oc.type);
                        DevBufP nb = make_buf(ctx, (size_t)(total > 0 ? total : 1) * w * words);
                        DxColumn dst = oc; dst.data = nb;
                        int rc = masks.assign(&dst, out_index); if (rc) return rc;
                        int64_t off = 0;
                        for (size_t k = 0; k < v.chunks.size(); k++) {
                            DxColumn sc = v.chunks[k]->cols[ci];
                            if (cs[k].n == 0) continue;
                            if (!sc.data) {        // this cached batch's constant (mixed): NULL rows stay NULL, a value is laid out for the batch's rows
                                if (sc.all_null || sc.type == DX_T_NULL) { off += cs[k].n; continue; }
                                int rc2 = materialize_const(&sc, v.chunks[k]->n_rows, mixed_dict); if (rc2) return rc2;
                            }
                            for (int wd = 0; wd < words; wd++) {
                                DxColumn view = sc;
                                if (oc.type == DX_T_FLAGS) {
                                    // word-major bitmap: gather each word plane as a 4-byte column
                                    view.type = DX_T_STR; view.vmask = nullptr;
                                    auto alias = std::make_shared<DevBuf>(); alias->p = sc.data->as<uint8_t>() + (size_t)wd * v.chunks[k]->n_rows * 4;
                                    view.data = alias;   // non-owning (ctx == nullptr => no release)
                                }
                                if (oc.type == DX_T_FLAGS) {
                                    int rc2 = dx_gather_column(ctx, view, cs[k].n, cs[k].filtered ? cs[k].sel->as<int64_t>() : nullptr,
                                                               nb->as<uint8_t>() + ((size_t)wd * total + off) * w, dst.vmask->as<uint32_t>() + off, dst.vbit);
                                    if (rc2) return rc2;
                                    continue;
                                }
                                // forwarded columns of one chunk that share a destination mask word go out together
                                PendingGather& pg = pending[k];
                                uint32_t* mdst = dst.vmask->as<uint32_t>() + off;
                                if (pg.jobs.n == DX_GATHER_MAX || (pg.jobs.n && pg.mask != mdst)) { int rc2 = flush_gather(k); if (rc2) return rc2; }
                                const int j = pg.jobs.n++;
                                pg.mask = mdst; pg.keep.push_back(view.data); pg.keep.push_back(view.vmask); pg.keep.push_back(nb); pg.keep.push_back(dst.vmask);
                                pg.jobs.in[j] = view.data->p; pg.jobs.vmask[j] = view.vmask ? view.vmask->p : nullptr; pg.jobs.out[j] = nb->as<uint8_t>() + (size_t)off * w;
                                pg.jobs.width[j] = (uint8_t)w; pg.jobs.vwidth[j] = (uint8_t)view.vwidth; pg.jobs.vbit_in[j] = (uint8_t)view.vbit; pg.jobs.vbit[j] = (uint8_t)dst.vbit;
                            }
                            off += cs[k].n;
                        }
                        oc = dst;
                    }
                    t->cols.push_back(oc); out_index++;
                }
                continue;
            }
            if (is_literal(e)) { t->cols.push_back(literal_column(e, item_name(it))); out_index++; continue; }
            // rule arrays -> flag bitmap column
            std::vector<ExprP> conds; std::shared_ptr<RuleTable> rt;
            if (match_rules(e, &conds, &rt)) {
                DxColumn oc; oc.name = item_name(it); oc.type = DX_T_FLAGS; oc.flag_words = ((int)conds.size() + 31) / 32; oc.aux = rt;
                oc.data = make_buf(ctx, (size_t)(total > 0 ? total : 1) * 4 * oc.flag_words);
                int64_t off = 0;
                for (size_t k = 0; k < v.chunks.size(); k++) {
                    if (cs[k].n == 0) continue;
                    if (v.chunks.size() > 1) return fail(DX_ERR_UNSUPPORTED, "rule evaluation over a multi-batch window view");
                    BindScope sc = scope_for(v.chunks[k].get(), core);
                    const int64_t* selp = cs[k].filtered ? cs[k].sel->as<int64_t>() : nullptr;
                    bool fast = false;
                    int rc = F->generic_rules_only ? DX_OK : dx_rules_eval(ctx, sc, conds, cs[k].n, selp, oc.data->as<uint32_t>() + off, oc.flag_words, &fast, &F->rule_plans[stmt_name + "#" + oc.name], F->rules_mode);
                    if (rc) return rc;
                    if (!fast) {   // rule set outside the bucketed planner's limits: generic interpreter
                        ExprPrograms P;
                        for (auto& c : conds) { rc = dx_compile_expr(ctx, sc, c, &P); if (rc) return rc; }
                        DevPrograms D; rc = dx_upload_programs(ctx, &P, &D); if (rc) return rc;
                        rc = dx_eval_flags(ctx, D, cs[k].n, selp, oc.data->as<uint32_t>() + off, oc.flag_words);
                        if (rc) return rc;
                    }
                    off += cs[k].n;
                }
                t->cols.push_back(oc); out_index++; continue;
            }
            // string-valued results (CONCAT, CAST AS STRING, user functions) and stringToTimestamp
            {
                BindScope psc = scope_for(proto, core); psc.qualifiers.push_back("Raw");
                if (needs_string_engine(e, psc)) {
                    DxColumn oc; oc.name = item_name(it);
                    int64_t off = 0; bool first = true; int w = 4;
                    for (size_t k = 0; k < v.chunks.size(); k++) {
                        BindScope sc = scope_for(v.chunks[k].get(), core); sc.qualifiers.push_back("Raw");
                        DxColumn full; int rc = eval_string_full(sc, v.chunks[k].get(), e, stmt_name + "#" + oc.name, &full); if (rc) return rc;
                        if (first) {
                            oc.type = full.type; oc.dict = full.dict; w = dx_type_width(full.type);
                            oc.data = make_buf(ctx, (size_t)(total > 0 ? total : 1) * w);
                            rc = masks.assign(&oc, out_index); if (rc) return rc;
                            first = false;
                        }
                        if (cs[k].n == 0 || full.all_null) { off += cs[k].n; continue; }
                        rc = dx_gather_column(ctx, full, cs[k].n, cs[k].filtered ? cs[k].sel->as<int64_t>() : nullptr,
                                              oc.data->as<uint8_t>() + (size_t)off * w, oc.vmask->as<uint32_t>() + off, oc.vbit);
                        if (rc) return rc;
                        off += cs[k].n;
                    }
                    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                    t->cols.push_back(oc); out_index++; continue;
                }
            }
            // computed column
            {
                DxColumn oc; oc.name = item_name(it);
                int rtype = 0; DevBufP data; int64_t off = 0;
                for (size_t k = 0; k < v.chunks.size(); k++) {
                    ExprPrograms P; BindScope sc = scope_for(v.chunks[k].get(), core); sc.qualifiers.push_back("Raw");
                    int rc = compile_scalar(sc, e, &P); if (rc) return rc;
                    if (k == 0) {
                        rtype = P.result_type[0]; if (rtype == DX_T_NULL) rtype = DX_T_I64;
                        data = make_buf(ctx, (size_t)(total > 0 ? total : 1) * dx_type_width(rtype));
                        oc.type = rtype; oc.data = data; oc.dict = P.result_dict[0];
                        rc = masks.assign(&oc, out_index); if (rc) return rc;
                    }
                    if (cs[k].n == 0) continue;
                    DevPrograms D; rc = dx_upload_programs(ctx, &P, &D); if (rc) return rc;
                    rc = dx_eval_column(ctx, D, 0, cs[k].n, cs[k].filtered ? cs[k].sel->as<int64_t>() : nullptr,
                                        data->as<uint8_t>() + (size_t)off * dx_type_width(rtype), dx_type_width(rtype), oc.vmask->as<uint32_t>() + off, oc.vbit);
                    if (rc) return rc;
                    off += cs[k].n;
                }
                t->cols.push_back(oc); out_index++;
            }
        }
        for (size_t k = 0; k < pending.size(); k++) { int rc = flush_gather(k); if (rc) return rc; }
        *out = t; return DX_OK;
    }

    // ---------------- string-valued expressions (dictionary columns) ----------------
    bool is_udf(const ExprP& e) const {
        if (e->kind != EK::Func) return false;
        auto it = F->udfs.find(e->sval);
        return it != F->udfs.end() && (int)e->args.size() == it->second.n_args;
    }
    bool string_valued(const ExprP& e, const BindScope& sc) const {
        if (e->kind == EK::Str) return true;
        if (e->kind == EK::Col) { std::string er; int ci = dx_resolve_column(sc, e->parts, &er); return ci >= 0 && sc.table->cols[ci].type == DX_T_STR; }
        if (e->kind == EK::Cast) return e->sval == "string" || e->sval == "varchar";
        if (e->kind == EK::Func) return e->sval == "CONCAT" || is_udf(e);
        return false;
    }
    bool needs_string_engine(const ExprP& e, const BindScope& sc) const {
        if (e->kind == EK::Func && e->sval == "STRINGTOTIMESTAMP" && e->args.size() == 1) return true;
        if (e->kind == EK::Cast && (e->sval == "string" || e->sval == "varchar")) return true;
        if (e->kind == EK::Func && (e->sval == "CONCAT" || is_udf(e))) return true;
        return false;
    }
    DxDictP result_dict(const std::string& key) {
        auto it = F->expr_dicts.find(key);
        if (it != F->expr_dicts.end()) return it->second;
        auto d = std::make_shared<DxDict>();
        if (d->init(ctx, 1u << 16, 4u << 20)) return nullptr;
        F->expr_dicts[key] = d;
        return d;
    }
    // Evaluates a string-engine expression over ALL rows of chunk t (row-aligned result).
    int eval_string_full(const BindScope& sc, const DxTable* t, const ExprP& e, const std::string& key, DxColumn* out) {
        const int64_t n = t->n_rows;
        if (e->kind == EK::Col) {
            std::string er; int ci = dx_resolve_column(sc, e->parts, &er);
            if (ci < 0) return fail(DX_ERR_ARG, er);
            *out = t->cols[ci]; return DX_OK;
        }
        DevBufP mask = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 4);
        DX_CUDA(ctx, cudaMemsetAsync(mask->p, 0, (size_t)(n > 0 ? n : 1) * 4, ctx->stream));
        out->vmask = mask; out->vwidth = 4; out->vbit = 0; out->is_const = false; out->all_null = false;
        if (e->kind == EK::Func && (e->sval == "STRINGTOTIMESTAMP" || is_udf(e))) {
            const bool is_ts = e->sval == "STRINGTOTIMESTAMP";
            const dx_flow::Udf u = is_ts ? dx_flow::Udf() : F->udfs[e->sval];
            // one text per call: the argument itself, or the arguments joined by a separator the callback wrapper splits at
            // (every distinct combination becomes one dictionary entry, so the host function runs once per combination)
            static const char SEP = '\x1f';
            auto call = [&](const std::string& joined, std::string* result, bool* is_null, std::string* err) {
                std::vector<char> buf(4096); int64_t ol = -1; int rcc;
                if (u.fnn) {
                    std::vector<std::string> parts; size_t a0 = 0;
                    for (int k = 0; k < u.n_args; k++) { size_t p1 = k + 1 < u.n_args ? joined.find(SEP, a0) : std::string::npos; parts.push_back(joined.substr(a0, p1 == std::string::npos ? std::string::npos : p1 - a0)); a0 = p1 == std::string::npos ? joined.size() : p1 + 1; }
                    std::vector<const char*> ptrs; std::vector<int64_t> lens;
                    for (auto& p2 : parts) { ptrs.push_back(p2.data()); lens.push_back((int64_t)p2.size()); }
                    rcc = u.fnn(u.user, u.n_args, ptrs.data(), lens.data(), buf.data(), (int64_t)buf.size(), &ol);
                } else rcc = u.fn(u.user, joined.data(), (int64_t)joined.size(), buf.data(), (int64_t)buf.size(), &ol);
                if (rcc != 0) { *err = "user function " + e->sval + " failed"; return; }
                *is_null = ol < 0; if (ol >= 0) result->assign(buf.data(), (size_t)ol);
            };
            if (!is_ts && u.n_args == 0) {     // Generator0 / UDF0: one call per batch, a constant column
                std::string res, uerr; bool is_null = true;
                call("", &res, &is_null, &uerr);
                if (!uerr.empty()) return fail(DX_ERR_STATE, uerr);
                out->vmask = nullptr; out->data = nullptr; out->is_const = true; out->type = DX_T_STR;
                if (is_null) { out->all_null = true; return DX_OK; }
                out->const_s = res; out->const_i = -1;
                return materialize_const(out, n, result_dict(key));          // callers of the string engine take row-aligned columns
            }
            ExprP arg_expr = e->args[0];
            if (!is_ts && u.n_args >= 2) {
                auto cc = std::make_shared<Expr>(); cc->kind = EK::Func; cc->sval = "CONCAT";
                for (int k = 0; k < u.n_args; k++) {
                    if (k) { auto sp = std::make_shared<Expr>(); sp->kind = EK::Str; sp->sval = std::string(1, SEP); cc->args.push_back(sp); }
                    cc->args.push_back(e->args[k]);
                }
                arg_expr = cc;
            }
            DxColumn arg; int rc = eval_string_full(sc, t, arg_expr, key + "/arg", &arg); if (rc) return rc;
            if (arg.all_null || arg.type == DX_T_NULL) { out->type = DX_T_STR; out->all_null = true; out->data = nullptr; return DX_OK; }
            if (arg.type != DX_T_STR) return fail(DX_ERR_UNSUPPORTED, e->sval + " needs a string argument");
            if (!arg.data) {   // constant argument: materialise through a one-entry dictionary
                rc = materialize_const(&arg, n, nullptr); if (rc) return rc;
            }
            char kb[64]; snprintf(kb, sizeof kb, "@%p", (void*)arg.dict.get());
            std::shared_ptr<LutCache>& lut = F->luts[to_string(e) + kb];
            if (!lut) { prune_luts(); lut = std::make_shared<LutCache>(); lut->src_dict = arg.dict; }   // the key is the dictionary's address: keep it alive
            if (is_ts) {
                rc = dx_lut_refresh(ctx, lut.get(), arg.dict, [](const std::string& s, long long* v) { return dx_string_to_timestamp_us(s, v); });
            } else {
                if (!lut->out_dict) { lut->out_dict = std::make_shared<DxDict>(); rc = lut->out_dict->init(ctx, 1u << 16, 4u << 20); if (rc) return rc; }
                DxDictP od = lut->out_dict; dx_ctx* c = ctx; std::string uerr;
                rc = dx_lut_refresh(ctx, lut.get(), arg.dict, [&](const std::string& s, long long* v) {
                    std::string res; bool is_null = true;
                    call(s, &res, &is_null, &uerr);
                    if (!uerr.empty() || is_null) return false;
                    int32_t code; if (od->insert_host(res, &code)) { uerr = c->last_error; return false; }
                    *v = code; return true;
                });
                if (!uerr.empty()) return fail(DX_ERR_STATE, uerr);
            }
            if (rc) return rc;
            out->type = is_ts ? DX_T_TS : DX_T_STR; out->dict = is_ts ? nullptr : lut->out_dict;
            const int w = is_ts ? 8 : 4;
            out->data = make_buf(ctx, (size_t)(n > 0 ? n : 1) * w);
            return dx_lut_apply(ctx, arg, *lut, n, nullptr, out->data->p, w, mask->as<uint32_t>(), 0);
        }
        // CONCAT / CAST AS STRING -> per-row byte assembly
        std::vector<ExprP> args;
        if (e->kind == EK::Cast) args.push_back(e->args[0]); else args = e->args;
        if (args.size() > DX_STRBUILD_MAX_PARTS) return fail(DX_ERR_UNSUPPORTED, "CONCAT with more than 16 arguments");
        StrBuildArgs a; memset(&a, 0, sizeof a);
        std::string lit; std::vector<DxColumn> keep;
        for (size_t k = 0; k < args.size(); k++) {
            StrPart& sp = a.parts[a.n_parts++];
            const ExprP& x = args[k];
            if (x->kind == EK::Str) { sp.kind = SP_LIT; sp.off = (uint32_t)lit.size(); sp.len = (uint32_t)x->sval.size(); lit += x->sval; continue; }
            if (x->kind == EK::Int) { std::string sv = std::to_string(x->ival); sp.kind = SP_LIT; sp.off = (uint32_t)lit.size(); sp.len = (uint32_t)sv.size(); lit += sv; continue; }
            DxColumn c;
            if (string_valued(x, sc)) { int rc = eval_string_full(sc, t, x, key + "/" + std::to_string(k), &c); if (rc) return rc; }
            else if (x->kind == EK::Col) {
                std::string er; int ci = dx_resolve_column(sc, x->parts, &er);
                if (ci < 0) return fail(DX_ERR_ARG, er);
                c = t->cols[ci];
            } else {   // numeric sub-expression: evaluate it first
                ExprPrograms P; int rc = dx_compile_expr(ctx, sc, x, &P); if (rc) return rc;
                DevPrograms D; rc = dx_upload_programs(ctx, &P, &D); if (rc) return rc;
                c.type = P.result_type[0]; c.data = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 8);
                c.vmask = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 4); c.vwidth = 4; c.vbit = 0;
                DX_CUDA(ctx, cudaMemsetAsync(c.vmask->p, 0, (size_t)(n > 0 ? n : 1) * 4, ctx->stream));
                rc = dx_eval_column(ctx, D, 0, n, nullptr, c.data->p, 8, c.vmask->as<uint32_t>(), 0); if (rc) return rc;
            }
            if (c.all_null || c.type == DX_T_NULL) { out->type = DX_T_STR; out->all_null = true; out->data = nullptr; return DX_OK; }
            if (!c.data) {   // constant column: its text
                std::string sv = c.type == DX_T_STR ? c.const_s : std::to_string(c.const_i);
                if (c.type == DX_T_F64) return fail(DX_ERR_UNSUPPORTED, "double to string formatting");
                sp.kind = SP_LIT; sp.off = (uint32_t)lit.size(); sp.len = (uint32_t)sv.size(); lit += sv; continue;
            }
            if (c.type == DX_T_F64 || c.type == DX_T_FLAGS) return fail(DX_ERR_UNSUPPORTED, "only strings, integers and timestamps can be turned into text here");
            sp.kind = c.type == DX_T_STR ? SP_STRCOL : (c.type == DX_T_TS ? SP_TS : SP_I64);
            sp.col.data = c.data->p; sp.col.vmask = c.vmask ? c.vmask->p : nullptr; sp.col.vwidth = c.vwidth; sp.col.vbit = c.vbit; sp.col.width = dx_type_width(c.type);
            if (c.type == DX_T_STR) { if (!c.dict) return fail(DX_ERR_STATE, "string column without dictionary"); sp.dict = c.dict->dev; }
            keep.push_back(c);
        }
        DxDictP od = result_dict(key);
        if (!od) return fail(DX_ERR_CUDA, "dictionary allocation failed");
        out->type = DX_T_STR; out->dict = od;
        out->data = make_buf(ctx, (size_t)(n > 0 ? n : 1) * 4);
        a.n_rows = n; a.sel = nullptr;
        return dx_strbuild(ctx, a, lit, od, out->data->as<int32_t>(), mask->as<uint32_t>(), 0);
    }

    // scalar expression with the two host-assisted built-ins handled here
    int compile_scalar(const BindScope& sc, const ExprP& e, ExprPrograms* P) {
        return dx_compile_expr(ctx, sc, e, P);
    }

    // ---------------- GROUP BY / DISTINCT / aggregates ----------------
    struct KeySpec { ExprP expr; int src_col = -1; bool is_const = false; DxColumn const_col; };
    struct AggSpec { std::string func; ExprP arg; bool star = false; int val_index = -1; int acc = -1, acc_cnt = -1; int arg_type = 0; };

    // COUNT(DISTINCT x) [with other decomposable aggregates]: two levels of the ordinary hash aggregation —
    // first per (keys, x), then per keys counting the non-NULL x of the distinct pairs.
    int exec_count_distinct(const SelectCore& core, const View& v, DxTableP* out) {
        ExprP dx; 
        for (auto& it : core.items) if (is_agg_func(it.expr) && it.expr->distinct) {
            if (it.expr->sval != "COUNT" || it.expr->args.size() != 1) return fail(DX_ERR_UNSUPPORTED, "only COUNT(DISTINCT x) is supported");
            if (dx && to_string(dx) != to_string(it.expr->args[0])) return fail(DX_ERR_UNSUPPORTED, "COUNT(DISTINCT) over two different expressions");
            dx = it.expr->args[0];
        }
        auto col = [](const std::string& n) { auto e = std::make_shared<Expr>(); e->kind = EK::Col; e->parts = {n}; return e; };
        auto agg = [](const std::string& f, ExprP a) { auto e = std::make_shared<Expr>(); e->kind = EK::Func; e->sval = f; if (a) e->args = {a}; else e->star_arg = true; return e; };
        SelectCore inner; inner.from = core.from; inner.from_alias = core.from_alias; inner.where = core.where;
        SelectCore outer; outer.from = "__distinct";
        for (size_t k = 0; k < core.group_by.size(); k++) {
            std::string nm = "__k" + std::to_string(k);
            inner.items.push_back({core.group_by[k], nm}); inner.group_by.push_back(core.group_by[k]);
            outer.group_by.push_back(col(nm));
        }
        inner.items.push_back({dx, "__dx"}); inner.group_by.push_back(dx);
        int na = 0;
        for (auto& it : core.items) {
            SelectItem oi; oi.alias = item_name(it);
            const ExprP& e = it.expr;
            if (is_agg_func(e)) {
                if (e->distinct) oi.expr = agg("COUNT", col("__dx"));
                else {
                    std::string nm = "__a" + std::to_string(na++);
                    const std::string& f = e->sval;
                    if (f == "AVG") return fail(DX_ERR_UNSUPPORTED, "AVG next to COUNT(DISTINCT)");
                    inner.items.push_back({e, nm});
                    oi.expr = agg((f == "COUNT") ? "SUM" : f, col(nm));
                }
            } else {
                int hit = -1;
                for (size_t k = 0; k < core.group_by.size(); k++) if (to_string(core.group_by[k]) == to_string(e)) hit = (int)k;
                if (hit >= 0) oi.expr = col("__k" + std::to_string(hit));
                else if (is_literal(e)) oi.expr = e;
                else return fail(DX_ERR_ARG, "'" + to_string(e) + "' is neither aggregated nor in GROUP BY");
            }
            outer.items.push_back(oi);
        }
        DxTableP mid; int rc = exec_agg(inner, v, &mid); if (rc) return rc;
        View mv; mv.chunks.push_back(mid);
        const std::string saved = stmt_name; stmt_name = saved + "#distinct";
        rc = exec_agg(outer, mv, out);
        stmt_name = saved;
        return rc;
    }

    int exec_agg(const SelectCore& core, const View& v, DxTableP* out) {
        for (auto& it : core.items) if (is_agg_func(it.expr) && it.expr->distinct) return exec_count_distinct(core, v, out);
        const DxTable* proto = v.chunks[0].get();
        BindScope psc = scope_for(proto, core);
        // expand the select list ('*' only makes sense with DISTINCT)
        std::vector<SelectItem> items;
        for (auto& it : core.items) {
            if (it.expr->kind == EK::Star) {
                for (size_t c = 0; c < proto->cols.size(); c++) {
                    SelectItem s; s.expr = std::make_shared<Expr>(); s.expr->kind = EK::Col; s.expr->parts = {proto->cols[c].name}; s.alias = proto->cols[c].name;
                    items.push_back(s);
                }
            } else items.push_back(it);
        }
        // group expressions: source columns first, then select aliases (spark.sql.groupByAliases)
        std::vector<ExprP> gexprs;
        if (!core.group_by.empty()) {
            for (auto& g : core.group_by) {
                ExprP ge = g;
                if (g->kind == EK::Col) {
                    std::string err; int ci = dx_resolve_column(psc, g->parts, &err);
                    if (ci < 0 && g->parts.size() == 1) {
                        for (auto& it : items) if (item_name(it) == g->parts[0] && !contains_agg(it.expr)) { ge = it.expr; break; }
                    }
                }
                gexprs.push_back(ge);
            }
        } else if (core.distinct) {
            for (auto& it : items) { if (contains_agg(it.expr)) return fail(DX_ERR_UNSUPPORTED, "DISTINCT with aggregates"); gexprs.push_back(it.expr); }
        }
        std::vector<KeySpec> keys;
        for (auto& ge : gexprs) {
            KeySpec k; k.expr = ge;
            if (is_literal(ge)) { k.is_const = true; k.const_col = literal_column(ge, ""); }
            else if (ge->kind == EK::Col) {
                std::string err; int ci = dx_resolve_column(psc, ge->parts, &err);
                if (ci < 0) return fail(DX_ERR_ARG, err);
                const DxColumn& c = proto->cols[ci];
                if (c.type == DX_T_FLAGS) return fail(DX_ERR_UNSUPPORTED, "grouping by a rule-flag column");
                if (!c.data || c.all_null) { k.is_const = true; k.const_col = c; } else k.src_col = ci;
            }
            keys.push_back(k);
        }
        // aggregates and the mapping of every select item
        std::vector<AggSpec> aggs; std::vector<ExprP> val_exprs; std::vector<AggAccDesc> accs;
        struct ItemMap { int key = -1; int agg = -1; bool literal = false; bool folded = false; };
        std::vector<ItemMap> imap(items.size());
        auto add_acc = [&](int kind, int col) { for (size_t a = 0; a < accs.size(); a++) if (accs[a].kind == kind && accs[a].col == col) return (int)a; accs.push_back({kind, col}); return (int)accs.size() - 1; };
        for (size_t i = 0; i < items.size(); i++) {
            const ExprP& e = items[i].expr;
            if (is_agg_func(e)) {
                AggSpec a; a.func = e->sval;
                if (e->distinct) return fail(DX_ERR_UNSUPPORTED, "COUNT(DISTINCT ...) is not yet inside the supported subset");
                if (e->star_arg || (a.func == "COUNT" && e->args.size() == 1 && is_literal(e->args[0]) && e->args[0]->kind != EK::Null)) { a.star = true; a.acc = add_acc(ACC_COUNT_STAR, -1); }
                else {
                    if (e->args.size() != 1) return fail(DX_ERR_UNSUPPORTED, a.func + " needs one argument");
                    a.arg = e->args[0];
                    int vi = -1; std::string key = to_string(a.arg);
                    for (size_t k2 = 0; k2 < val_exprs.size(); k2++) if (to_string(val_exprs[k2]) == key) vi = (int)k2;
                    if (vi < 0) { val_exprs.push_back(a.arg); vi = (int)val_exprs.size() - 1; }
                    a.val_index = vi;
                }
                aggs.push_back(a); imap[i].agg = (int)aggs.size() - 1;
            } else if (contains_agg(e)) return fail(DX_ERR_UNSUPPORTED, "expressions over aggregates: '" + to_string(e) + "'");
            else {
                std::string key = to_string(e); int hit = -1;
                for (size_t k2 = 0; k2 < keys.size(); k2++) if (to_string(keys[k2].expr) == key) hit = (int)k2;
                if (hit < 0 && e->kind == EK::Col) {   // same column under another spelling
                    std::string err; int ci = dx_resolve_column(psc, e->parts, &err);
                    for (size_t k2 = 0; k2 < keys.size(); k2++) if (ci >= 0 && keys[k2].src_col == ci) hit = (int)k2;
                }
                if (hit >= 0) imap[i].key = hit;
                else if (is_literal(e)) imap[i].literal = true;
                else if (!has_column_ref(e)) imap[i].folded = true;     // foldable: DATE_TRUNC('second', current_timestamp()) ...
                else return fail(DX_ERR_ARG, "'" + key + "' is neither aggregated nor in GROUP BY");
            }
        }
        if ((int)val_exprs.size() > DX_AGG_MAX_ACCS) return fail(DX_ERR_UNSUPPORTED, "too many aggregate arguments");
        // type the aggregate arguments against the first chunk
        std::vector<int> val_type(val_exprs.size()), val_src(val_exprs.size(), -1);
        std::vector<DxDictP> val_dict(val_exprs.size());
        for (size_t k2 = 0; k2 < val_exprs.size(); k2++) {
            if (val_exprs[k2]->kind == EK::Col) {
                std::string err; int ci = dx_resolve_column(psc, val_exprs[k2]->parts, &err);
                if (ci < 0) return fail(DX_ERR_ARG, err);
                if (proto->cols[ci].data && !proto->cols[ci].all_null) { val_src[k2] = ci; val_type[k2] = proto->cols[ci].type; val_dict[k2] = proto->cols[ci].dict; continue; }
            }
            ExprPrograms P; int rc = dx_compile_expr(ctx, psc, val_exprs[k2], &P); if (rc) return rc;
            val_type[k2] = P.result_type[0] == DX_T_NULL ? DX_T_I64 : P.result_type[0];
        }
        for (auto& a : aggs) {
            if (a.star) continue;
            const int vt = val_type[a.val_index]; a.arg_type = vt;
            const bool f = vt == DX_T_F64;
            a.acc_cnt = add_acc(ACC_COUNT, a.val_index);
            if (a.func == "COUNT") a.acc = a.acc_cnt;
            else if (vt == DX_T_STR) return fail(DX_ERR_UNSUPPORTED, a.func + " over a string column");
            else if (a.func == "SUM") a.acc = add_acc(f ? ACC_SUM_F : ACC_SUM_I, a.val_index);
            else if (a.func == "AVG") {
                if (f) a.acc = add_acc(ACC_SUM_F, a.val_index);
                else {   // AVG(long) accumulates CAST(x AS DOUBLE): a derived double value column
                    auto cast = std::make_shared<Expr>(); cast->kind = EK::Cast; cast->sval = "double"; cast->args = {val_exprs[a.val_index]};
                    val_exprs.push_back(cast); val_type.push_back(DX_T_F64); val_src.push_back(-1); val_dict.push_back(nullptr);
                    a.acc = add_acc(ACC_SUM_F, (int)val_exprs.size() - 1);
                }
            }
            else if (a.func == "MIN") a.acc = add_acc(f ? ACC_MIN_F : ACC_MIN_I, a.val_index);
            else if (a.func == "MAX") a.acc = add_acc(f ? ACC_MAX_F : ACC_MAX_I, a.val_index);
        }
        if (accs.empty()) accs.push_back({ACC_COUNT_STAR, -1});
        std::vector<int> hashed;   // keys that take part in hashing
        for (size_t k2 = 0; k2 < keys.size(); k2++) if (!keys[k2].is_const) hashed.push_back((int)k2);
        if ((int)hashed.size() > DX_AGG_MAX_KEYS) return fail(DX_ERR_UNSUPPORTED, "more than 16 group-by keys");

        // selections per chunk
        struct ChunkSel { DevBufP sel; int64_t n = 0; bool filtered = false; };
        std::vector<ChunkSel> cs(v.chunks.size()); int64_t total = 0;
        const bool fuse_range = v.ranged && !core.where;
        for (size_t k = 0; k < v.chunks.size(); k++) {
            if (fuse_range) {
                // no WHERE: the window's time range is applied inside the aggregation kernel, no selection pass
                const DxTableP& ch = v.chunks[k];
                const bool outside = ch->has_ts_bounds && (ch->ts_max < v.lo_us || ch->ts_min >= v.hi_us);
                cs[k].filtered = false; cs[k].n = outside ? 0 : ch->n_rows;
            } else {
                int rc = chunk_selection(v, v.chunks[k], core, &cs[k].sel, &cs[k].n, &cs[k].filtered); if (rc) return rc;
            }
            total += cs[k].n;
        }
        unsigned long long cap = 1024;
        auto hint = F->agg_capacity_hint.find(stmt_name);
        {   // remembered size of this statement's table, but never far below what this batch's row count could need
            unsigned long long want = (unsigned long long)std::min<int64_t>(total, 1 << 22) * 2;
            if (hint != F->agg_capacity_hint.end()) { cap = hint->second; want = std::min<unsigned long long>(want, 1ull << 20); }
            while (cap < want) cap <<= 1;
        }
        AggTable T; unsigned long long n_groups = 0;
        std::vector<int> key_type(keys.size(), DX_T_I64); std::vector<DxDictP> key_dict(keys.size());
        // One chunk's columns as the aggregation kernel's input; `hold` keeps computed columns alive until the kernel has run.
        struct InputHold { std::vector<DevBufP> temps; DevBufP tmask; std::vector<DxColumn> string_keys; };
        const std::vector<ExprP>* input_vals = &val_exprs;     // which value columns make_input lays out (a shared pane partial may carry more)
        auto make_input = [&](const DxTable* ch, const long long* sel, int64_t n, bool with_range, AggInput* inp, InputHold* hold) -> int {
            AggInput& in = *inp; memset(&in, 0, sizeof in);
            BindScope sc = scope_for(ch, core);
            in.sel = sel; in.n_rows = n;
            if (with_range) {
                std::string err; int tc = dx_resolve_column(sc, std::vector<std::string>{v.ts_col}, &err);
                if (tc < 0) return fail(DX_ERR_ARG, err);
                const DxColumn& c = ch->cols[tc];
                if (!c.data || c.type != DX_T_TS) return fail(DX_ERR_UNSUPPORTED, "window timestamp column must be a materialised timestamp");
                in.range_col.data = c.data->p; in.range_col.vmask = c.vmask ? c.vmask->p : nullptr; in.range_col.vwidth = c.vwidth;
                in.range_col.vbit = c.vbit; in.range_col.width = 8;
                in.range_lo = ch->has_cut ? std::max(v.lo_us, ch->ts_cut) : v.lo_us; in.range_hi = v.hi_us; in.has_range = 1;
            }
            auto col_desc = [&](const DxColumn& c) { ProgCol pc; pc.data = c.data->p; pc.vmask = c.vmask ? c.vmask->p : nullptr; pc.vwidth = c.vwidth; pc.vbit = c.vbit; pc.width = dx_type_width(c.type); pc.pad = 0; return pc; };
            auto eval_temp = [&](const ExprP& e, int* type_out, ProgCol* pc) -> int {
                ExprPrograms P; int rc2 = dx_compile_expr(ctx, sc, e, &P); if (rc2) return rc2;
                DevPrograms D; rc2 = dx_upload_programs(ctx, &P, &D); if (rc2) return rc2;
                int ty = P.result_type[0] == DX_T_NULL ? DX_T_I64 : P.result_type[0];
                DevBufP b = make_buf(ctx, (size_t)ch->n_rows * dx_type_width(ty)); hold->temps.push_back(b);
                if (!hold->tmask) { hold->tmask = make_buf(ctx, (size_t)ch->n_rows * 4); DX_CUDA(ctx, cudaMemsetAsync(hold->tmask->p, 0, (size_t)ch->n_rows * 4, ctx->stream)); }
                int bit = (int)hold->temps.size() - 1;
                if (bit >= 32) return fail(DX_ERR_UNSUPPORTED, "too many computed group/aggregate expressions");
                rc2 = dx_eval_column(ctx, D, 0, ch->n_rows, nullptr, b->p, dx_type_width(ty), hold->tmask->as<uint32_t>(), bit); if (rc2) return rc2;
                pc->data = b->p; pc->vmask = hold->tmask->p; pc->vwidth = 4; pc->vbit = bit; pc->width = dx_type_width(ty); pc->pad = 0;
                *type_out = ty; return DX_OK;
            };
            int rc = DX_OK;
            for (size_t hk = 0; hk < hashed.size(); hk++) {
                KeySpec& ks = keys[hashed[hk]];
                if (ks.expr->kind == EK::Col) {
                    std::string err; int ci = dx_resolve_column(sc, ks.expr->parts, &err);
                    if (ci < 0) return fail(DX_ERR_ARG, err);
                    const DxColumn& c = ch->cols[ci];
                    if (!c.data) return fail(DX_ERR_UNSUPPORTED, "group key constant in one batch and materialised in another");
                    in.keys[hk] = col_desc(c); key_type[hashed[hk]] = c.type;
                    if (c.type == DX_T_STR) {
                        if (key_dict[hashed[hk]] && key_dict[hashed[hk]] != c.dict) return fail(DX_ERR_UNSUPPORTED, "group key dictionaries differ between batches");
                        key_dict[hashed[hk]] = c.dict;
                    }
                } else if (needs_string_engine(ks.expr, sc)) {
                    // CONCAT(...) / CAST AS STRING / user function as a group key (e.g. SELECT DISTINCT ... CONCAT(...) AS Pivot1)
                    DxColumn full; rc = eval_string_full(sc, ch, ks.expr, stmt_name + "#key" + std::to_string(hashed[hk]), &full); if (rc) return rc;
                    if (!full.data || full.all_null) return fail(DX_ERR_UNSUPPORTED, "constant string expression as a group key");
                    hold->string_keys.push_back(full);     // keeps the buffers alive until the kernel has run
                    in.keys[hk] = col_desc(full); key_type[hashed[hk]] = full.type;
                    if (key_dict[hashed[hk]] && key_dict[hashed[hk]] != full.dict) return fail(DX_ERR_UNSUPPORTED, "group key dictionaries differ between batches");
                    key_dict[hashed[hk]] = full.dict;
                } else { int ty; rc = eval_temp(ks.expr, &ty, &in.keys[hk]); if (rc) return rc; key_type[hashed[hk]] = ty; }
            }
            const std::vector<ExprP>& ivals = *input_vals;
            for (size_t vi = 0; vi < ivals.size(); vi++) {
                bool done = false;
                if (ivals[vi]->kind == EK::Col) {
                    std::string err; int ci = dx_resolve_column(sc, ivals[vi]->parts, &err);
                    if (ci >= 0 && ch->cols[ci].data && !ch->cols[ci].all_null) { in.vals[vi] = col_desc(ch->cols[ci]); done = true; }
                }
                if (!done) { int ty; rc = eval_temp(ivals[vi], &ty, &in.vals[vi]); if (rc) return rc; }
            }
            return DX_OK;
        };

        // Pane partials: a window over several cached batches whose ends fall on batch-interval boundaries is answered from one
        // pre-aggregated table per cached batch (built once, when the batch is first seen by this statement).
        const long long pane_us = v.pane_us;
        bool plain_keys = true;
        for (int hk : hashed) if (keys[hk].expr->kind != EK::Col) plain_keys = false;
        const bool pane_mode = fuse_range && !F->pane_partials_off && pane_us > 0 && v.chunks.size() >= 2 && !stmt_name.empty() && plain_keys &&
                               v.lo_us % pane_us == 0 && v.hi_us % pane_us == 0 && (int)hashed.size() + 1 <= DX_AGG_MAX_KEYS;
        // column arrays filled directly by the shared-memory path (otherwise extracted from T below)
        std::vector<DevBufP> pre_k, pre_a; DevBufP pre_knull; bool have_cols = false;
        const bool pane_smem = pane_mode && !F->pane_smem_off && dx_aggp_fits((int)hashed.size(), (int)accs.size());
        if (pane_smem) {
            // Partials are shared between statements: a statement takes any partial family member ("variant") of the same
            // window source, pane width and group keys whose accumulators cover its own (Tumbling's COUNT/SUM/MIN/MAX partial
            // also answers a 5-minute COUNT/SUM), so a cached batch is scattered and pre-aggregated once, not once per statement.
            std::string fam = "P|" + std::to_string(pane_us) + "|" + v.ts_col;
            for (int hk : hashed) fam += "|k:" + to_string(keys[hk].expr);
            std::vector<std::string> need(accs.size());
            for (size_t a = 0; a < accs.size(); a++) need[a] = std::to_string(accs[a].kind) + ":" + (accs[a].col >= 0 ? to_string(val_exprs[accs[a].col]) : std::string("*"));
            auto& variants = F->pane_variants[fam];
            auto covers = [](const std::vector<std::string>& have, const std::vector<std::string>& want) {
                for (auto& w : want) if (std::find(have.begin(), have.end(), w) == have.end()) return false;
                return true;
            };
            int vi = -1;
            for (size_t q = 0; q < variants.size(); q++)
                if (covers(variants[q].desc, need) && (vi < 0 || variants[q].desc.size() < variants[vi].desc.size())) vi = (int)q;
            if (vi < 0) {
                // a variant the new one covers is superseded: its statements find the wider one from their next batch on
                for (size_t q = 0; q < variants.size();) {
                    if (covers(need, variants[q].desc)) { F->pane_partials_p.erase(variants[q].sig); F->pane_P.erase(variants[q].sig); variants.erase(variants.begin() + q); }
                    else q++;
                }
                dx_flow::PaneVariant nv2; nv2.vals = val_exprs; nv2.accs = accs; nv2.desc = need;
                nv2.sig = fam + "|#" + std::to_string(F->pane_variant_seq++);
                variants.push_back(nv2); vi = (int)variants.size() - 1;
            }
            const dx_flow::PaneVariant var = variants[vi];       // by value: the vector may be edited by a later statement
            std::vector<int> amap(accs.size());
            for (size_t a = 0; a < accs.size(); a++) amap[a] = (int)(std::find(var.desc.begin(), var.desc.end(), need[a]) - var.desc.begin());
            const std::string sig = var.sig;
            auto& cache = F->pane_partials_p[sig];
            const int nk = (int)hashed.size();
            for (auto it2 = cache.begin(); it2 != cache.end();) {
                bool live = false; for (auto& chp : v.chunks) if (chp.get() == it2->first) live = true;
                if (!live) it2 = cache.erase(it2); else ++it2;
            }
            int& P = F->pane_P[sig];
            for (int round = 0;; round++) {
                if (round > 6) return fail(DX_ERR_CAPACITY, "pane partitions overflow");
                bool again = false;
                for (size_t k = 0; k < v.chunks.size() && !again; k++) {
                    if (cs[k].n == 0) continue;
                    const DxTableP& chp = v.chunks[k];
                    if (cache.count(chp.get())) {       // only the key types / dictionaries of this batch's columns are needed
                        BindScope sc = scope_for(chp.get(), core);
                        for (size_t hk = 0; hk < hashed.size(); hk++) {
                            std::string err; int ci = dx_resolve_column(sc, keys[hashed[hk]].expr->parts, &err);
                            if (ci < 0) return fail(DX_ERR_ARG, err);
                            key_type[hashed[hk]] = chp->cols[ci].type;
                            if (chp->cols[ci].type == DX_T_STR) {
                                if (key_dict[hashed[hk]] && key_dict[hashed[hk]] != chp->cols[ci].dict) return fail(DX_ERR_UNSUPPORTED, "group key dictionaries differ between batches");
                                key_dict[hashed[hk]] = chp->cols[ci].dict;
                            }
                        }
                        continue;
                    }
                    InputHold hold; AggInput in;
                    input_vals = &var.vals;            // the partial carries the variant's value columns, not only this statement's
                    int rc = make_input(chp.get(), nullptr, chp->n_rows, true, &in, &hold);
                    input_vals = &val_exprs;
                    if (rc) return rc;
                    if (P == 0 && F->pane_partitions_first > 0) P = F->pane_partitions_first;      // tests start too coarse on purpose
                    if (P == 0) { P = 256; while (P < 16384 && (long long)P * 1200 < chp->n_rows) P <<= 1; }
                    in.has_range = chp->has_cut ? 1 : 0; in.range_lo = chp->ts_cut; in.range_hi = 0x7FFFFFFFFFFFFFFFll; in.pane_us = pane_us;   // every row of the cached batch, keyed by its pane
                    PanePartialP pp; pp.chunk = chp;
                    bool ovf = false;
                    auto& fill = F->pane_fill[sig];
                    rc = dx_aggp_build(ctx, in, nk, (int)var.vals.size(), var.accs, P, &pp.ap, &ovf, &fill.first, &fill.second); if (rc) return rc;
                    if (ovf) { again = true; break; }
                    ctx->counters["k3_partial_rows_built"] += (long long)pp.ap.n;
                    cache[chp.get()] = std::move(pp);
                }
                if (!again) {
                    std::vector<const AggPPartial*> parts; long long rows = 0;
                    for (size_t k = 0; k < v.chunks.size(); k++) if (cs[k].n) { const AggPPartial* ap = &cache[v.chunks[k].get()].ap; parts.push_back(ap); rows += ap->n; }
                    long long maxg = rows;
                    auto hint2 = F->agg_capacity_hint.find(stmt_name);
                    if (hint2 != F->agg_capacity_hint.end()) maxg = std::min<long long>(rows, (long long)hint2->second);
                    for (int attempt = 0; attempt < 2; attempt++) {
                        const size_t nr = (size_t)(maxg > 0 ? maxg : 1);
                        pre_k.assign(nk, nullptr); pre_a.assign(accs.size(), nullptr);
                        long long* kp[DX_AGG_MAX_KEYS]; unsigned long long* ap2[DX_AGG_MAX_ACCS];
                        for (int k2 = 0; k2 < nk; k2++) { pre_k[k2] = make_buf(ctx, nr * 8); kp[k2] = pre_k[k2]->as<long long>(); }
                        for (size_t a2 = 0; a2 < accs.size(); a2++) { pre_a[a2] = make_buf(ctx, nr * 8); ap2[a2] = pre_a[a2]->as<unsigned long long>(); }
                        pre_knull = make_buf(ctx, nr * 8);
                        long long ng = 0; bool ovf = false;
                        int rc = dx_aggp_combine(ctx, parts, nk, accs, amap, v.lo_us / pane_us, v.hi_us / pane_us, maxg, kp, pre_knull->as<unsigned long long>(), ap2, &ng, &ovf);
                        if (rc) return rc;
                        if (!ovf) { n_groups = (unsigned long long)ng; have_cols = true; break; }
                        if (maxg < rows) { ctx->counters["k3p_output_regrows"]++; maxg = rows; continue; }     // the remembered size was too small: the row count always suffices
                        again = true; break;                              // a partition overflowed its shared table
                    }
                }
                ctx->counters["k3p_partitions"] = P;
                if (!again) break;
                ctx->counters["k3p_repartitions"]++;
                P = std::min(P * 4, 1 << 16); cache.clear(); F->pane_fill.erase(sig);   // finer partitions; every partial of this signature is rebuilt
                if (P >= (1 << 16) && round > 3) return fail(DX_ERR_CAPACITY, "pane partitions overflow");
            }
        } else
        if (pane_mode) {
            std::string sig = std::to_string(pane_us) + "|" + v.ts_col;     // the views of one flow range over the same cached batches
            for (int hk : hashed) sig += "|k:" + to_string(keys[hk].expr);
            for (auto& ve : val_exprs) sig += "|v:" + to_string(ve);
            for (auto& ac : accs) sig += "|a:" + std::to_string(ac.kind) + "," + std::to_string(ac.col);
            auto& cache = F->pane_partials[sig];      // statements with the same keys, values and accumulators share the partials
            const int nk = (int)hashed.size();
            // evict partials of batches that left the window cache
            for (auto it2 = cache.begin(); it2 != cache.end();) {
                bool live = false; for (auto& chp : v.chunks) if (chp.get() == it2->first) live = true;
                if (!live || it2->second.sig != sig) it2 = cache.erase(it2); else ++it2;
            }
            for (size_t k = 0; k < v.chunks.size(); k++) {
                if (cs[k].n == 0) continue;
                const DxTableP& chp = v.chunks[k];
                if (cache.count(chp.get())) {       // still fills key_type / key_dict for this batch's columns
                    BindScope sc = scope_for(chp.get(), core);
                    for (size_t hk = 0; hk < hashed.size(); hk++) {
                        KeySpec& ks = keys[hashed[hk]];
                        if (ks.expr->kind != EK::Col) continue;
                        std::string err; int ci = dx_resolve_column(sc, ks.expr->parts, &err);
                        if (ci < 0) return fail(DX_ERR_ARG, err);
                        key_type[hashed[hk]] = chp->cols[ci].type;
                        if (chp->cols[ci].type == DX_T_STR) {
                            if (key_dict[hashed[hk]] && key_dict[hashed[hk]] != chp->cols[ci].dict) return fail(DX_ERR_UNSUPPORTED, "group key dictionaries differ between batches");
                            key_dict[hashed[hk]] = chp->cols[ci].dict;
                        }
                    }
                    continue;
                }
                PanePartial pp; pp.chunk = chp; pp.sig = sig;
                // sized from the previous batch's partial (groups are stable from batch to batch), else from the row count
                unsigned long long pcap = 1024, pwant = (unsigned long long)std::min<int64_t>(chp->n_rows, 1 << 22) * 2;   // first guess; an overflow retries 4x larger
                auto phint = F->agg_capacity_hint.find("#pane|" + sig);
                if (phint != F->agg_capacity_hint.end()) pwant = std::min(pwant, phint->second);
                while (pcap < pwant) pcap <<= 1;
                for (int attempt = 0;; attempt++) {
                    AggTable P; int rc = P.init(ctx, nk + 1, (int)val_exprs.size(), accs, pcap); if (rc) return rc;
                    InputHold hold; AggInput in;
                    rc = make_input(chp.get(), nullptr, chp->n_rows, true, &in, &hold); if (rc) return rc;
                    in.has_range = chp->has_cut ? 1 : 0; in.range_lo = chp->ts_cut; in.range_hi = 0x7FFFFFFFFFFFFFFFll; in.pane_us = pane_us;          // every row of the cached batch, keyed by its pane
                    memset(&in.keys[nk], 0, sizeof(ProgCol));
                    rc = P.update(in); if (rc) return rc;
                    unsigned long long ng = 0; bool ovf = false;
                    rc = P.finish(&ng, &ovf); if (rc) return rc;
                    if (ovf) { if (attempt > 12) return fail(DX_ERR_CAPACITY, "pane partial table overflow"); pcap = P.capacity * 4; continue; }
                    pp.n = (int64_t)ng;
                    const size_t nr = (size_t)(ng > 0 ? ng : 1);
                    long long* kp[DX_AGG_MAX_KEYS]; unsigned long long* ap[DX_AGG_MAX_ACCS];
                    for (int k2 = 0; k2 <= nk; k2++) { pp.keys.push_back(make_buf(ctx, nr * 8)); kp[k2] = pp.keys.back()->as<long long>(); }
                    for (size_t a2 = 0; a2 < accs.size(); a2++) { pp.accs.push_back(make_buf(ctx, nr * 8)); ap[a2] = pp.accs.back()->as<unsigned long long>(); }
                    DevBufP knull = make_buf(ctx, nr * 8); pp.valid32 = make_buf(ctx, nr * 4);
                    rc = P.extract(ng, kp, knull->as<unsigned long long>(), ap); if (rc) return rc;
                    rc = dx_agg_keynull_to_valid(ctx, knull->as<unsigned long long>(), (long long)ng, pp.valid32->as<uint32_t>()); if (rc) return rc;
                    DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                    ctx->counters["k3_partial_rows_built"] += (long long)ng;
                    { unsigned long long h2 = 1024; while (h2 < ng * 2 + 16) h2 <<= 1; F->agg_capacity_hint["#pane|" + sig] = h2 * 2; }
                    break;
                }
                cache[chp.get()] = std::move(pp);
            }
            // combine: COUNTs become sums of partial counts, everything else keeps its operator
            std::vector<AggAccDesc> accs2(accs.size());
            for (size_t a2 = 0; a2 < accs.size(); a2++) {
                accs2[a2].kind = (accs[a2].kind == ACC_COUNT_STAR || accs[a2].kind == ACC_COUNT) ? ACC_SUM_I : accs[a2].kind;
                accs2[a2].col = (int)a2;
            }
            for (int attempt = 0;; attempt++) {
                int rc = T.init(ctx, nk, (int)accs.size(), accs2, cap); if (rc) return rc;
                for (size_t k = 0; k < v.chunks.size(); k++) {
                    if (cs[k].n == 0) continue;
                    const PanePartial& pp = cache[v.chunks[k].get()];
                    if (pp.n == 0) continue;
                    AggInput in; memset(&in, 0, sizeof in);
                    in.n_rows = pp.n;
                    for (int hk = 0; hk < nk; hk++) { ProgCol pc; pc.data = pp.keys[hk]->p; pc.vmask = pp.valid32->p; pc.vwidth = 4; pc.vbit = hk; pc.width = 8; pc.pad = 0; in.keys[hk] = pc; }
                    for (size_t a2 = 0; a2 < accs.size(); a2++) { ProgCol pc; pc.data = pp.accs[a2]->p; pc.vmask = nullptr; pc.vwidth = 0; pc.vbit = 0; pc.width = 8; pc.pad = 0; in.vals[a2] = pc; }
                    in.range_col.data = pp.keys[nk]->p; in.range_col.vmask = nullptr; in.range_col.width = 8;
                    in.range_lo = v.lo_us / pane_us; in.range_hi = v.hi_us / pane_us; in.has_range = 1;
                    rc = T.update(in); if (rc) return rc;
                }
                bool overflow = false;
                rc = T.finish(&n_groups, &overflow); if (rc) return rc;
                if (!overflow) break;
                if (attempt > 12) return fail(DX_ERR_CAPACITY, "aggregation table overflow");
                cap = T.capacity * 4;
            }
        } else
        for (int attempt = 0;; attempt++) {
            int rc = T.init(ctx, (int)hashed.size(), (int)val_exprs.size(), accs, cap); if (rc) return rc;
            for (size_t k = 0; k < v.chunks.size(); k++) {
                if (cs[k].n == 0) continue;
                InputHold hold; AggInput in;
                rc = make_input(v.chunks[k].get(), cs[k].filtered ? cs[k].sel->as<long long>() : nullptr, cs[k].n, fuse_range, &in, &hold); if (rc) return rc;
                rc = T.update(in); if (rc) return rc;
                DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // computed columns die with this scope
            }
            bool overflow = false;
            int rc2 = T.finish(&n_groups, &overflow); if (rc2) return rc2;
            if (!overflow) break;
            if (attempt > 12) return fail(DX_ERR_CAPACITY, "aggregation table overflow");
            cap = T.capacity * 4;
        }
        if (!stmt_name.empty()) { unsigned long long h = 1024; while (h < n_groups * 2 + 16) h <<= 1; F->agg_capacity_hint[stmt_name] = have_cols ? h : std::max(h, std::min(T.capacity, h * 2)); }

        // global aggregate over zero rows -> one row (count 0, others NULL)
        const bool global = core.group_by.empty() && !core.distinct;
        int64_t n_out = (int64_t)n_groups;
        const bool synth = global && n_groups == 0;
        if (synth) n_out = 1;
        // pull the groups out
        std::vector<DevBufP> kbufs(hashed.size()), abufs(accs.size());
        DevBufP knull = make_buf(ctx, (size_t)(n_out > 0 ? n_out : 1) * 8);
        long long* kptr[DX_AGG_MAX_KEYS]; unsigned long long* aptr[DX_AGG_MAX_ACCS];
        if (have_cols && !synth) {
            kbufs = pre_k; abufs = pre_a; knull = pre_knull;
            for (size_t k = 0; k < hashed.size(); k++) kptr[k] = kbufs[k]->as<long long>();
            for (size_t a = 0; a < accs.size(); a++) aptr[a] = abufs[a]->as<unsigned long long>();
        } else {
        for (size_t k = 0; k < hashed.size(); k++) { kbufs[k] = make_buf(ctx, (size_t)(n_out > 0 ? n_out : 1) * 8); kptr[k] = kbufs[k]->as<long long>(); }
        for (size_t a = 0; a < accs.size(); a++) { abufs[a] = make_buf(ctx, (size_t)(n_out > 0 ? n_out : 1) * 8); aptr[a] = abufs[a]->as<unsigned long long>(); }
        }
        if (have_cols && !synth) {
        } else if (synth) {
            DX_CUDA(ctx, cudaMemsetAsync(knull->p, 0, 8, ctx->stream));
            for (auto& b : abufs) DX_CUDA(ctx, cudaMemsetAsync(b->p, 0, 8, ctx->stream));
        } else { int rc = T.extract(n_groups, kptr, knull->as<unsigned long long>(), aptr); if (rc) return rc; }

        auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = n_out;
        MaskAlloc masks(ctx, n_out);
        FinJobs fin; fin.n = 0; uint32_t* fin_mask = nullptr; std::vector<DevBufP> fin_keep;
        auto flush_fin = [&]() {
            if (!fin.n) return;
            k_finalize_multi<<<(unsigned)std::min<long long>((n_out + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, ctx->stream>>>(fin, knull->as<unsigned long long>(), n_out, fin_mask);
            ctx->launches["k_finalize"]++;
            fin.n = 0; fin_mask = nullptr;
        };
        auto finalize = [&](int mode, const unsigned long long* a, const unsigned long long* cnt, int keybit, DxColumn* oc, int index) -> int {
            int w = dx_type_width(oc->type);
            oc->data = make_buf(ctx, (size_t)(n_out > 0 ? n_out : 1) * w);
            int rc = masks.assign(oc, index); if (rc) return rc;
            if (n_out > 0) {
                uint32_t* m = oc->vmask->as<uint32_t>();
                if (fin.n == FIN_MAX || (fin.n && fin_mask != m)) flush_fin();
                const int j = fin.n++;
                fin_mask = m; fin_keep.push_back(oc->data); fin_keep.push_back(oc->vmask);
                fin.mode[j] = mode; fin.keybit[j] = keybit; fin.width[j] = w; fin.vbit[j] = oc->vbit; fin.a[j] = a; fin.cnt[j] = cnt; fin.out[j] = oc->data->p;
            }
            return DX_OK;
        };
        for (size_t i = 0; i < items.size(); i++) {
            DxColumn oc; oc.name = item_name(items[i]);
            if (imap[i].literal) { oc = literal_column(items[i].expr, oc.name); t->cols.push_back(oc); continue; }
            if (imap[i].folded) {
                ExprPrograms P; int rc = dx_compile_expr(ctx, psc, items[i].expr, &P); if (rc) return rc;
                DevPrograms D; rc = dx_upload_programs(ctx, &P, &D); if (rc) return rc;
                oc.type = P.result_type[0] == DX_T_NULL ? DX_T_I64 : P.result_type[0];
                oc.data = make_buf(ctx, (size_t)(n_out > 0 ? n_out : 1) * 8);
                rc = masks.assign(&oc, (int)i); if (rc) return rc;
                rc = dx_eval_column(ctx, D, 0, n_out, nullptr, oc.data->p, 8, oc.vmask->as<uint32_t>(), oc.vbit); if (rc) return rc;
                DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                t->cols.push_back(oc); continue;
            }
            if (imap[i].key >= 0) {
                KeySpec& ks = keys[imap[i].key];
                if (ks.is_const) { DxColumn c = ks.const_col; c.name = oc.name; t->cols.push_back(c); continue; }
                int hk = (int)(std::find(hashed.begin(), hashed.end(), imap[i].key) - hashed.begin());
                oc.type = key_type[imap[i].key]; oc.dict = key_dict[imap[i].key];
                int rc = finalize(0, reinterpret_cast<unsigned long long*>(kptr[hk]), nullptr, hk, &oc, (int)i); if (rc) return rc;
                t->cols.push_back(oc); continue;
            }
            const AggSpec& a = aggs[imap[i].agg];
            int rc;
            if (a.func == "COUNT") { oc.type = DX_T_I64; rc = finalize(1, aptr[a.acc], nullptr, 0, &oc, (int)i); }
            else if (a.func == "AVG") { oc.type = DX_T_F64; rc = finalize(3, aptr[a.acc], aptr[a.acc_cnt], 0, &oc, (int)i); }
            else { oc.type = (a.func == "SUM" && a.arg_type != DX_T_F64) ? DX_T_I64 : a.arg_type; if (a.func != "SUM" && a.arg_type == DX_T_TS) oc.type = DX_T_TS; rc = finalize(2, aptr[a.acc], aptr[a.acc_cnt], 0, &oc, (int)i); }
            if (rc) return rc;
            t->cols.push_back(oc);
        }
        flush_fin();
        // DISTINCT / GROUP BY over MAP/STRUCT leaves keeps their shaping
        for (size_t i = 0; i < items.size() && i < t->cols.size(); i++) {
            DxColumn& oc = t->cols[i]; oc.nest_path.clear(); oc.nest_kinds.clear();
            if (items[i].expr->kind != EK::Col) continue;
            std::string err; int ci = dx_resolve_column(psc, items[i].expr->parts, &err);
            if (ci >= 0 && proto->cols[ci].name == oc.name) { oc.nest_path = proto->cols[ci].nest_path; oc.nest_kinds = proto->cols[ci].nest_kinds; }
        }
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        *out = t; return DX_OK;
    }
};

// =======================================================================================
// flow construction
// =======================================================================================
static bool parse_state_schema(const std::string& s, std::vector<std::pair<std::string, int>>* out) {
    std::stringstream ss(s); std::string part;
    while (std::getline(ss, part, ',')) {
        std::string p = trim(part); if (p.empty()) continue;
        size_t sp = p.find_first_of(" \t:");
        if (sp == std::string::npos) return false;
        std::string name = trim(p.substr(0, sp)), ty = lower(trim(p.substr(sp + 1)));
        if (!ty.empty() && ty[0] == ':') ty = trim(ty.substr(1));
        int t = (ty == "long" || ty == "bigint" || ty == "int" || ty == "integer") ? DX_T_I64 : (ty == "double" || ty == "float") ? DX_T_F64 :
                (ty == "string") ? DX_T_STR : (ty == "timestamp") ? DX_T_TS : (ty == "boolean") ? DX_T_BOOL : 0;
        if (!t) return false;
        out->emplace_back(name, t);
    }
    return !out->empty();
}

static DxTableP empty_table(dx_ctx* ctx, const std::vector<std::pair<std::string, int>>& schema) {
    auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = 0;
    for (auto& kv : schema) { DxColumn c; c.name = kv.first; c.type = kv.second; c.data = make_buf(ctx, 8); t->cols.push_back(c); }
    return t;
}

int dx_flow_build(dx_ctx* ctx, const char* conf_text, const char* schema_json, const char* const* projection_steps, int n_steps,
                  const char* transform_text, dx_flow** out) {
    std::unique_ptr<dx_flow> F(new dx_flow()); F->ctx = ctx;
    parse_conf(conf_text ? conf_text : "", &F->conf);
    F->decoder.reset(new dx_decoder());
    int rc = F->decoder->init(ctx, strip_bom(schema_json)); if (rc) return rc;
    // projection steps: one selectExpr string per line (ProjectionHandler.scala:23-36)
    for (int s = 0; s < n_steps; s++) {
        std::vector<SelectItem> items;
        for (auto& line : split_lines(strip_bom(projection_steps[s]))) {
            std::string l = trim(line); if (l.empty()) continue;
            try { Parser p(l); items.push_back(p.select_item()); }
            catch (SqlError& e) { return ctx->set_error(DX_ERR_PARSE, "projection '" + l + "': " + e.what()); }
        }
        if (!items.empty()) F->projections.push_back(items);
    }
    // transform
    std::string err;
    rc = dx_parse_transform(transform_text ? transform_text : "", &F->transform, &err);
    if (rc) return ctx->set_error(rc, err);
    for (auto& c : F->transform.commands) {
        if (!c.is_query) return ctx->set_error(DX_ERR_UNSUPPORTED, "non-query statement in transform: '" + c.text.substr(0, 80) + "'");
        try { Parser p(c.text); c.stmt = p.statement(); }
        catch (SqlError& e) { return ctx->set_error(DX_ERR_PARSE, "statement '" + c.name + "': " + e.what()); }
    }
    // time windows (TimeWindowHandler.scala:41-67)
    const std::string root = "datax.job.process.";
    for (auto& kv : F->conf) {
        const std::string pre = root + "timewindow.";
        if (kv.first.compare(0, pre.size(), pre) == 0) {
            std::string rest = kv.first.substr(pre.size());
            size_t dot = rest.rfind('.');
            if (dot != std::string::npos && rest.substr(dot + 1) == "windowduration") {
                bool ok; long long s = dx_parse_duration_seconds(kv.second, &ok);
                if (!ok) return ctx->set_error(DX_ERR_PARSE, "bad window duration '" + kv.second + "'");
                F->windows.emplace_back(rest.substr(0, dot), s);
                F->max_window_s = std::max(F->max_window_s, s);
            }
        }
        const std::string spre = root + "statetable.";
        if (kv.first.compare(0, spre.size(), spre) == 0) {
            std::string rest = kv.first.substr(spre.size());
            size_t dot = rest.rfind('.');
            if (dot != std::string::npos && rest.substr(dot + 1) == "schema") {
                StateTable st; st.name = rest.substr(0, dot);
                if (!parse_state_schema(kv.second, &st.schema)) return ctx->set_error(DX_ERR_PARSE, "bad state table schema '" + kv.second + "'");
                st.phys[0] = empty_table(ctx, st.schema); st.phys[1] = empty_table(ctx, st.schema);
                F->state_tables[st.name] = st;
            }
        }
        const std::string opre = "datax.job.output.";
        if (kv.first.compare(0, opre.size(), opre) == 0) {
            std::string rest = kv.first.substr(opre.size());
            size_t dot = rest.find('.');
            std::string name = dot == std::string::npos ? rest : rest.substr(0, dot);
            if (std::find(F->outputs.begin(), F->outputs.end(), name) == F->outputs.end()) F->outputs.push_back(name);
        }
    }
    for (auto& kv : F->conf)      // PropertiesHandler.buildAppendProperties: datax.job.process.appendproperty.<key>=<value>
        if (kv.first.compare(0, (root + "appendproperty.").size(), root + "appendproperty.") == 0)
            F->append_properties.emplace_back(kv.first.substr((root + "appendproperty.").size()), kv.second);
    { auto g = F->conf.find(root + "gpu.genericrules"); F->generic_rules_only = g != F->conf.end() && lower(g->second) == "true"; }
    { auto g = F->conf.find(root + "gpu.rulesstaging"); if (g != F->conf.end() && lower(g->second) == "false") F->rules_mode |= DX_RULES_NO_STAGING; }
    { auto g = F->conf.find(root + "gpu.rulesfast"); if (g != F->conf.end() && lower(g->second) == "false") F->rules_mode |= DX_RULES_NO_FAST; }
    { auto g = F->conf.find(root + "gpu.panepartials"); F->pane_partials_off = g != F->conf.end() && lower(g->second) == "false"; }
    { auto g = F->conf.find(root + "gpu.panesmem"); F->pane_smem_off = g != F->conf.end() && lower(g->second) == "false"; }
    { auto g = F->conf.find(root + "gpu.panepartitions"); if (g != F->conf.end()) { int v = atoi(g->second.c_str()); int p2 = 1; while (p2 < v) p2 <<= 1; F->pane_partitions_first = v > 0 ? p2 : 0; } }
    auto wm = F->conf.find(root + "watermark"); auto tc = F->conf.find(root + "timestampcolumn");
    F->windows_enabled = !F->windows.empty() && wm != F->conf.end() && tc != F->conf.end();
    if (F->windows_enabled) {
        bool ok; F->watermark_s = dx_parse_duration_seconds(wm->second, &ok);
        if (!ok) return ctx->set_error(DX_ERR_PARSE, "bad watermark '" + wm->second + "'");
        F->ts_column = tc->second;
    } else { F->max_window_s = 0; }
    *out = F.release();
    return DX_OK;
}

// Reference data: CSV text -> a small device table whose columns are all strings, as
// CSVUtil.loadCSVReferenceData does (datax-utility CSVUtil.scala:27-29: header option, no schema inference).
int dx_flow_add_refdata_impl(dx_flow* F, const std::string& name, const std::string& csv, char delim, bool header) {
    dx_ctx* ctx = F->ctx;
    std::vector<std::vector<std::string>> rows; std::vector<std::vector<bool>> nulls;
    {
        std::string text = strip_bom(csv);
        std::vector<std::string> cur; std::vector<bool> curn; std::string field; bool inq = false, quoted = false; bool any = false;
        auto end_field = [&]() { cur.push_back(field); curn.push_back(field.empty() && !quoted); field.clear(); quoted = false; };
        auto end_row = [&]() { end_field(); if (!(cur.size() == 1 && cur[0].empty() && !any)) { rows.push_back(cur); nulls.push_back(curn); } cur.clear(); curn.clear(); any = false; };
        for (size_t i = 0; i < text.size(); i++) {
            char c = text[i];
            if (inq) { if (c == '"') { if (i + 1 < text.size() && text[i + 1] == '"') { field.push_back('"'); i++; } else inq = false; } else field.push_back(c); continue; }
            if (c == '"') { inq = true; quoted = true; any = true; }
            else if (c == delim) { end_field(); any = true; }
            else if (c == '\n') { if (!field.empty() && field.back() == '\r') field.pop_back(); end_row(); }
            else { field.push_back(c); any = true; }
        }
        if (!field.empty() || !cur.empty()) { if (!field.empty() && field.back() == '\r') field.pop_back(); end_row(); }
    }
    if (rows.empty()) return ctx->set_error(DX_ERR_PARSE, "reference data '" + name + "' is empty");
    std::vector<std::string> names;
    size_t ncol = rows[0].size();
    if (header) { names = rows[0]; rows.erase(rows.begin()); nulls.erase(nulls.begin()); }
    else for (size_t c = 0; c < ncol; c++) names.push_back("_c" + std::to_string(c));
    auto t = std::make_shared<DxTable>(); t->ctx = ctx; t->n_rows = (int64_t)rows.size(); t->name = name;
    for (size_t c = 0; c < ncol; c++) {
        DxColumn col; col.name = names[c]; col.type = DX_T_STR;
        col.dict = std::make_shared<DxDict>();
        int rc = col.dict->init(ctx, 1u << 12, 1u << 20); if (rc) return rc;
        std::vector<int32_t> codes(rows.size(), -1); std::vector<uint32_t> mask(rows.size(), 0);
        for (size_t r = 0; r < rows.size(); r++) {
            if (c >= rows[r].size() || nulls[r][c]) continue;       // empty field -> NULL (Spark CSV nullValue default)
            int32_t code; rc = col.dict->insert_host(rows[r][c], &code); if (rc) return rc;
            codes[r] = code; mask[r] = 1;
        }
        col.data = make_buf(ctx, rows.size() * 4 + 16); col.vmask = make_buf(ctx, rows.size() * 4 + 16); col.vwidth = 4; col.vbit = 0;
        DX_CUDA(ctx, cudaMemcpyAsync(col.data->p, codes.data(), rows.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        DX_CUDA(ctx, cudaMemcpyAsync(col.vmask->p, mask.data(), rows.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        t->cols.push_back(col);
    }
    F->ref_tables[name] = t;
    return DX_OK;
}

// =======================================================================================
// one micro-batch: processDataset -> project -> route (CommonProcessorFactory.scala:333-399)
// =======================================================================================
static DxColumn meta_const(const std::string& name, bool has, const std::string& v, std::vector<std::string> path, const std::string& kinds) {
    DxColumn c; c.name = name; c.is_const = true; c.nest_path = std::move(path); c.nest_kinds = kinds;
    if (has) { c.type = DX_T_STR; c.const_s = v; c.const_i = -1; } else { c.type = DX_T_NULL; c.all_null = true; }
    return c;
}
// java.sql.Timestamp.toString in the UTC session: yyyy-mm-dd hh:mm:ss.f (at least one fractional digit)
static std::string java_timestamp_string(long long ms) {
    long long s = ms / 1000, frac = ms % 1000; if (frac < 0) { frac += 1000; s--; }
    time_t tt = (time_t)s; struct tm g; gmtime_r(&tt, &g);
    char b[64]; snprintf(b, sizeof b, "%04d-%02d-%02d %02d:%02d:%02d", g.tm_year + 1900, g.tm_mon + 1, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec);
    std::string out = b;
    if (frac == 0) return out + ".0";
    char f[8]; snprintf(f, sizeof f, "%03lld", frac); std::string fs = f; while (fs.size() > 1 && fs.back() == '0') fs.pop_back();
    return out + "." + fs;
}
void dx_flow::add_metadata_columns(DxTable* t, long long batch_ms, long long now_ms) const {
    const BatchInfo& b = batch_info;
    const std::string fi = "__DataX_FileInfo";
    t->cols.push_back(meta_const(fi + ".inputPath", b.has_input_path, b.input_path, {"inputPath"}, "s"));
    if (b.has_output_folders && !b.output_folders.empty())
        for (auto& kv : b.output_folders) t->cols.push_back(meta_const(fi + ".outputFolders." + kv.first, true, kv.second, {"outputFolders", kv.first}, "sm"));
    else t->cols.push_back(meta_const(fi + ".outputFolders", false, "", {"outputFolders"}, "s"));
    t->cols.push_back(meta_const(fi + ".outputFileName", b.has_output_file_name, b.output_file_name, {"outputFileName"}, "s"));
    {
        DxColumn c; c.name = fi + ".fileTime"; c.is_const = true; c.nest_path = {"fileTime"}; c.nest_kinds = "s";
        if (b.has_file_time) { c.type = DX_T_TS; c.const_i = b.file_time_ms * 1000; } else { c.type = DX_T_NULL; c.all_null = true; }
        t->cols.push_back(c);
    }
    t->cols.push_back(meta_const(fi + ".ruleIndexPrefix", b.has_rule_index_prefix, b.rule_index_prefix, {"ruleIndexPrefix"}, "s"));
    t->cols.push_back(meta_const(fi + ".target", b.has_target, b.target, {"target"}, "s"));
    // DataXProperties (PropertiesHandler.scala:26-35): appendproperty.* of the conf, then the processing properties; null values dropped
    std::vector<std::pair<std::string, std::string>> props = append_properties;
    auto put = [&](const std::string& k, const std::string& v) { for (auto& kv : props) if (kv.first == k) { kv.second = v; return; } props.emplace_back(k, v); };
    put("BatchTime", java_timestamp_string(batch_ms));
    if (b.has_file_time) put("BlobTime", java_timestamp_string(b.file_time_ms));
    if (b.has_output_file_name) put("BlobPathHint", b.output_file_name);
    put("CPTime", java_timestamp_string(now_ms));
    put("CPExecutor", b.executor_id);
    for (auto& kv : props) t->cols.push_back(meta_const("DataXProperties." + kv.first, true, kv.second, {kv.first}, "m"));
    for (auto& kv : b.properties) t->cols.push_back(meta_const("Properties." + kv.first, true, kv.second, {kv.first}, "m"));
    for (auto& kv : b.system_properties) t->cols.push_back(meta_const("SystemProperties." + kv.first, true, kv.second, {kv.first}, "m"));
}
int dx_flow::set_event_property(int which, const char* key, const uint8_t* bytes, const int64_t* offsets, const uint8_t* valid, int64_t n_events) {
    if ((which != 0 && which != 1) || !key || !*key || n_events < 0 || (n_events > 0 && (!bytes || !offsets)))
        return ctx->set_error(DX_ERR_ARG, "dx_flow_set_event_property: bad arguments");
    EventProperty ep; ep.column = std::string(which ? "SystemProperties." : "Properties.") + key;
    DxDictP& dict = event_prop_dicts[ep.column];
    if (!dict) { dict = std::make_shared<DxDict>(); int rc = dict->init(ctx, 1u << 16, 1u << 22); if (rc) return rc; }
    ep.codes.assign((size_t)n_events, -1); ep.valid.assign((size_t)n_events, 0u);
    std::map<std::string, int32_t> seen; std::vector<std::string> distinct; std::vector<int32_t> which_str((size_t)n_events, -1);
    for (int64_t r = 0; r < n_events; r++) {
        if (valid && !valid[r]) continue;
        if (offsets[r + 1] < offsets[r]) return ctx->set_error(DX_ERR_ARG, "dx_flow_set_event_property: offsets are not ascending");
        std::string sv(reinterpret_cast<const char*>(bytes) + offsets[r], (size_t)(offsets[r + 1] - offsets[r]));
        auto hit = seen.find(sv);
        if (hit == seen.end()) { hit = seen.emplace(sv, (int32_t)distinct.size()).first; distinct.push_back(sv); }
        which_str[(size_t)r] = hit->second; ep.valid[(size_t)r] = 1u;
    }
    std::vector<int32_t> codes;
    int rc = dict->insert_host_many(distinct, &codes); if (rc) return rc;
    for (int64_t r = 0; r < n_events; r++) if (which_str[(size_t)r] >= 0) ep.codes[(size_t)r] = codes[(size_t)which_str[(size_t)r]];
    for (auto& e : event_props) if (e.column == ep.column) { e = std::move(ep); return DX_OK; }
    event_props.push_back(std::move(ep));
    return DX_OK;
}
// the per-event map values become dictionary columns beside Raw, replacing a per-batch constant of the same key
int dx_flow::add_event_property_columns(DxTable* t) {
    for (auto& ep : event_props) {
        if ((int64_t)ep.codes.size() != t->n_rows)
            return ctx->set_error(DX_ERR_ARG, "event property '" + ep.column + "' was given for " + std::to_string(ep.codes.size()) + " events, the batch has " + std::to_string(t->n_rows));
        DxColumn c; c.name = ep.column; c.type = DX_T_STR; c.dict = event_prop_dicts[ep.column];
        c.nest_path = {ep.column.substr(ep.column.find('.') + 1)}; c.nest_kinds = "m";
        const size_t nr = (size_t)(t->n_rows > 0 ? t->n_rows : 1);
        c.data = make_buf(ctx, nr * 4); c.vmask = make_buf(ctx, nr * 4); c.vwidth = 4; c.vbit = 0;
        if (t->n_rows > 0) {
            DX_CUDA(ctx, cudaMemcpyAsync(c.data->p, ep.codes.data(), nr * 4, cudaMemcpyHostToDevice, ctx->stream));
            DX_CUDA(ctx, cudaMemcpyAsync(c.vmask->p, ep.valid.data(), nr * 4, cudaMemcpyHostToDevice, ctx->stream));
            DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        }
        bool replaced = false;
        for (auto& old : t->cols) if (old.name == c.name) { old = c; replaced = true; }
        if (!replaced) t->cols.push_back(c);
    }
    event_props.clear();
    return DX_OK;
}
bool dx_is_metadata_column(const std::string& n) {
    return n.compare(0, 8, "__DataX_") == 0 || n.compare(0, 16, "DataXProperties.") == 0 || n.compare(0, 11, "Properties.") == 0 || n.compare(0, 17, "SystemProperties.") == 0;
}

// project(): from_json + selectExpr steps (CommonProcessorFactory.scala:90-103)
int dx_flow::project(const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events, int memory_kind,
                     long long now_ms, DxTableP* out, long long batch_ms) {
    const long long batch_ms_for_props = batch_ms >= 0 ? batch_ms : now_ms;
    NvtxRange nvtx_project("project");
    DxTableP raw;
    int rc = memory_kind == 1 ? decoder->decode_device(json, n_bytes, offsets, n_events, &raw) : decoder->decode_host(json, n_bytes, offsets, n_events, &raw);
    if (rc) return rc;
    ViewMap views;
    Exec X{this, ctx, &views, now_ms * 1000, ""};
    // the frame the ingest adapters build around Raw (CPF:405-421, 458-461, 628-654): __DataX_FileInfo beside every event, and
    // what project() adds before the steps (CPF:97): DataXProperties; Properties / SystemProperties when the batch carries them
    DxTableP projected = std::make_shared<DxTable>(*raw);
    add_metadata_columns(projected.get(), batch_ms_for_props, now_ms);
    rc = add_event_property_columns(projected.get()); if (rc) return rc;
    bool first = true;
    for (auto& step : projections) {
        View v; v.chunks.push_back(projected); views["Raw"] = v;
        SelectCore core; core.from = "Raw"; core.items = step;
        X.stmt_name = "projection";
        DxTableP next; rc = X.exec_plain(core, views["Raw"], &next); if (rc) return rc;
        // selectExpr(projection ++ preservedColumns) (CPF:95-100): the __DataX_* columns survive every step
        for (auto& c : projected->cols) {
            if (c.name.compare(0, 8, "__DataX_") != 0) continue;
            bool have = false; for (auto& n : next->cols) if (n.name == c.name) have = true;
            if (!have) next->cols.push_back(c);
        }
        projected = next; first = false;
    }
    if (first) {       // no projection step: the frame as it is, minus what only project() sees
        auto keep = std::make_shared<DxTable>(*projected); keep->cols.clear();
        for (auto& c : projected->cols) if (!dx_is_metadata_column(c.name) || c.name.compare(0, 8, "__DataX_") == 0) keep->cols.push_back(c);
        projected = keep;
    }
    *out = projected;
    return DX_OK;
}

int dx_flow::process(const std::string& name_space, const uint8_t* json, int64_t n_bytes, const int64_t* offsets, int64_t n_events,
                     int memory_kind, long long batch_ms, long long interval_ms, long long now_ms, dx_result** out) {
    DxTableP projected;
    const auto t0 = std::chrono::steady_clock::now();
    batch_t0 = t0; batch_t0_set = true;
    if (on_interval) on_interval(on_interval_user, batch_ms);                 // CPF:351-353
    for (auto& kv : udfs) if (kv.second.refresh)                              // answers of a refreshing function are not kept across batches
        for (auto it = luts.begin(); it != luts.end();) { if (it->first.compare(0, kv.first.size() + 1, kv.first + "(") == 0) it = luts.erase(it); else ++it; }
    normalizer_seconds = 0;
    if (normalizer) {
        if (memory_kind == 1) return ctx->set_error(DX_ERR_UNSUPPORTED, "an input normalizer is set (user code on the host): the batch must be passed in host memory");
        if (offsets) return ctx->set_error(DX_ERR_UNSUPPORTED, "an input normalizer needs NDJSON framing (the normalised records have other offsets)");
        const auto n0 = std::chrono::steady_clock::now();
        normalized.resize((size_t)n_bytes * 2 + 64);
        int64_t n_out = -1;
        if (normalizer(normalizer_user, json, n_bytes, normalized.data(), (int64_t)normalized.size(), &n_out) != 0 || n_out < 0 || n_out > (int64_t)normalized.size())
            return ctx->set_error(DX_ERR_STATE, "the input normalizer failed");
        json = normalized.data(); n_bytes = n_out;
        normalizer_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - n0).count();
    }
    int rc = project(json, n_bytes, offsets, n_events, memory_kind, now_ms, &projected, batch_ms);
    if (rc) return rc;
    const auto t1 = std::chrono::steady_clock::now();
    pending_unparsed = (long long)decoder->last_stats.limited_rows;
    rc = route(name_space, projected, batch_ms, interval_ms, now_ms, out);
    pending_unparsed = 0; batch_t0_set = false;
    ctx->counters["host_us:project"] += std::chrono::duration_cast<std::chrono::microseconds>(t1 - t0).count();
    ctx->counters["host_us:route"] += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t1).count();
    return rc;
}

// route() on an already projected table (CommonProcessorFactory.scala:131-328).  Multi-GPU runs call this on the
// rows a rank received from the key-hash shuffle.
int dx_flow::route(const std::string& name_space, const DxTableP& projected, long long batch_ms, long long interval_ms, long long now_ms,
                   dx_result** out) {
    std::unique_ptr<dx_result> R(new dx_result()); R->flow = this;
    auto metric = [&](const std::string& k, double v) { R->metric_keys.push_back(k); R->metric_vals.push_back(v); };
    if (!batch_t0_set) batch_t0 = std::chrono::steady_clock::now();       // route() called on its own (dx_flow_process_table)
    NvtxRange nvtx_route("route");
    int rc = DX_OK;
    ViewMap views;
    Exec X{this, ctx, &views, now_ms * 1000, ""};
    // ---- route() (CPF:131-328)
    metric("Input_Normalized_Events_Count", (double)projected->n_rows);
    // events Spark would have parsed that this decoder made NULL rows (dx_decoder_limited_rows); the key appears only when there are any
    if (pending_unparsed > 0) metric("Input_Unparsed_Events_Count", (double)pending_unparsed);
    NamespaceState& NS = ns[name_space];
    if (windows_enabled) {
        const long long T = batch_ms;
        const long long E = T - watermark_s * 1000, S = E - max_window_s * 1000;
        DxTableP proj_w = projected;
        {   // e.g. `current_timestamp() AS eventTimeStamp` (HomeAutomationLocal.json:290): the window column is a per-batch
            // constant; give it a buffer so the cached rows carry their own timestamps
            BindScope bs; bs.table = projected.get();
            std::string err; int tc = dx_resolve_column(bs, std::vector<std::string>{ts_column}, &err);
            if (tc < 0) return ctx->set_error(DX_ERR_ARG, "timestamp column: " + err);
            if (!projected->cols[tc].data && !projected->cols[tc].all_null) {
                proj_w = std::make_shared<DxTable>(*projected);
                rc = X.materialize_const(&proj_w->cols[tc], proj_w->n_rows, nullptr); if (rc) return rc;
            }
        }
        // The late-event drop (CPF:161-167: rows with timestamp >= E stay) is kept as a predicate on the cached batch, not as a
        // filtered copy: cur shares the projected table's buffers, and every reader of a cached batch — window views, pane partial
        // builds — applies ts >= ts_cut together with its own time range.  One pass over the timestamp column gives the count of
        // the kept rows (the metric) and their timestamp bounds (later window views skip batches that cannot intersect them).
        DxTableP cur = std::make_shared<DxTable>(*proj_w);
        cur->has_cut = true; cur->ts_cut = E * 1000; cur->has_ts_bounds = false;
        long long kept = 0;
        {
            BindScope bs; bs.table = cur.get();
            std::string err; int tc = dx_resolve_column(bs, std::vector<std::string>{ts_column}, &err);
            if (tc < 0) return ctx->set_error(DX_ERR_ARG, "timestamp column: " + err);
            const DxColumn& c = cur->cols[tc];
            if (cur->n_rows > 0 && !c.all_null && c.type != DX_T_NULL) {
                if (!c.data || c.type != DX_T_TS) return ctx->set_error(DX_ERR_UNSUPPORTED, "window timestamp column must be a materialised timestamp");
                DevBuf mm(ctx, 32);
                long long init[3] = {0x7FFFFFFFFFFFFFFFll, (long long)0x8000000000000000ull, 0};
                DX_CUDA(ctx, cudaMemcpyAsync(mm.p, init, 24, cudaMemcpyHostToDevice, ctx->stream));
                k_minmax_i64<<<(unsigned)std::min<long long>((cur->n_rows + 255) / 256, (long long)ctx->sm_count * 8), 256, 0, ctx->stream>>>(
                    c.data->as<long long>(), c.vmask ? c.vmask->p : nullptr, c.vwidth, c.vbit, cur->n_rows, cur->ts_cut, mm.as<long long>());
                ctx->launches["k_minmax"]++;
                long long h[3];
                DX_CUDA(ctx, cudaMemcpyAsync(h, mm.p, 24, cudaMemcpyDeviceToHost, ctx->stream));
                DX_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                kept = h[2];
                if (h[2] > 0) { cur->has_ts_bounds = true; cur->ts_min = h[0]; cur->ts_max = h[1]; }
            }
            if (kept == 0) { cur->has_ts_bounds = true; cur->ts_min = 1; cur->ts_max = 0; }      // nothing kept: outside every window
        }
        metric("Input_DataXProcessedInput_Events_Count", (double)kept);
        const long long cut = T - (watermark_s + max_window_s) * 1000;
        for (auto it = NS.past_batches.begin(); it != NS.past_batches.end();) { if (it->first <= cut) it = NS.past_batches.erase(it); else ++it; }
        std::vector<DxTableP> U; U.push_back(cur);
        if (NS.past_batches.size() > 1) for (auto& kv : NS.past_batches) U.push_back(kv.second);   // the reference's size>1 quirk (CPF:198-201)
        auto ranged = [&](long long lo_ms, long long hi_ms) { View v; v.chunks = U; v.ranged = true; v.ts_col = ts_column; v.lo_us = lo_ms * 1000; v.hi_us = hi_ms * 1000; v.pane_us = interval_ms * 1000; return v; };
        views["DataXProcessedInput_Window"] = ranged(S, E);
        for (auto& w : windows) views[w.first] = ranged(E - w.second * 1000, E);
        views["DataXProcessedInput"] = ranged(E - (interval_ms / 1000) * 1000, E);
        View b; b.chunks.push_back(projected); views["DataXProcessedInput_Batch"] = b;
        NS.past_batches[T] = cur;
    } else {
        metric("Input_DataXProcessedInput_Events_Count", (double)projected->n_rows);
        View v; v.chunks.push_back(projected); views["DataXProcessedInput"] = v;
    }
    for (auto& kv : state_tables) { View v; v.chunks.push_back(kv.second.phys[kv.second.active]); views[kv.first] = v; }
    for (auto& kv : ref_tables) { View v; v.chunks.push_back(kv.second); views[kv.first] = v; }
    for (auto& cmd : transform.commands) {
        X.stmt_name = cmd.name;
        const auto t_stmt = std::chrono::steady_clock::now();
        DxTableP res;
        { NvtxRange nvtx_stmt("stmt:" + cmd.name); rc = X.exec_statement(cmd.stmt, &res); }
        if (rc) return rc;
        ctx->counters["host_us:stmt:" + cmd.name] += std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t_stmt).count();
        res->name = cmd.name;
        auto st = state_tables.find(cmd.name);
        if (st != state_tables.end()) {
            // INSERT OVERWRITE standby + flip (StateTableHandler.scala:99-115; CPF:259-264)
            StateTable& S2 = st->second;
            if (res->cols.size() != S2.schema.size()) return ctx->set_error(DX_ERR_ARG, "state table '" + cmd.name + "': query yields " + std::to_string(res->cols.size()) + " columns, schema has " + std::to_string(S2.schema.size()));
            for (size_t c = 0; c < res->cols.size(); c++) res->cols[c].name = S2.schema[c].first;   // INSERT is positional
            S2.phys[1 - S2.active] = res; S2.active = 1 - S2.active; S2.flipped = !S2.flipped;
            View v; v.chunks.push_back(res); views[cmd.name] = v;
        } else {
            View v; v.chunks.push_back(res); views[cmd.name] = v;
        }
        R->datasets[cmd.name] = res; R->dataset_order.push_back(cmd.name);
    }
    for (auto& o : outputs) {
        auto it = R->datasets.find(o);
        if (it == R->datasets.end()) return ctx->set_error(DX_ERR_ARG, "could not find data set name '" + o + "' for output '" + o + "'");
        // what OutputManager.sink reports before it fans out (OutputManager.scala:121-122: df.count() as Sink_InputEvents, prefixed
        // Output_<name>_ at CPF:304); the sinks' own counters (Sink_<sinker>_All ...) are added by the unchanged sinks on the host side
        metric("Output_" + o + "_Sink_InputEvents", (double)it->second->n_rows);
    }
    R->datasets["DataXProcessedInput_Batch"] = projected;
    // CPF:372-377: seconds spent on the batch, and the age of the batch when its processing ended (wall clock)
    if (normalizer) metric("Latency-Normalizer", normalizer_seconds);       // user code on the host, outside the GPU path: visible, not hidden
    metric("Latency-Process", std::chrono::duration<double>(std::chrono::steady_clock::now() - batch_t0).count());
    metric("Latency-Batch", (double)(std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::system_clock::now().time_since_epoch()).count() - batch_ms) / 1000.0);
    for (auto& k : R->metric_keys) R->metric_key_ptrs.push_back(k.c_str());
    *out = R.release();
    return DX_OK;
}
