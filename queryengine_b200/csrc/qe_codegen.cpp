// qe_codegen.cpp -- see qe_codegen.h
#include "qe_codegen.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <vector>

namespace qe {
namespace {

struct SV {                  // one entry of the compile-time operand stack
    std::string val;
    std::string valid;       // "true" when the value is statically non-null
    VType t = VT_NONE;
    bool is_const = false;
    double k = 0.0;
    std::string raw_int;     // non-empty: the value is a bare INT64 / DATE32 column, this is its integer payload
    int lit = -1;            // is_const: slot of the literal in QeParams::lit_* (-1: the value is embedded in the code)
};

struct Gen {
    std::string body;
    int ntmp = 0;
    // common subexpressions: (function, operand names) -> the temporary that already holds the result.
    // The reference evaluates a repeated subtree every time it meets it (it has no DAG); the functions
    // are pure, so evaluating once is indistinguishable -- and a shared subtree with a divide or a MOD
    // in it is most of the work of a row.
    std::map<std::string, SV> memo;

    std::string tmp(const char *prefix)
    {
        char b[32];
        snprintf(b, sizeof b, "%s%d", prefix, ntmp++);
        return b;
    }
    void line(const std::string &s) { body += "        " + s + "\n"; }
};

std::string f64_literal(double k)
{
    uint64_t bits;
    memcpy(&bits, &k, sizeof bits);
    char b[64];
    snprintf(b, sizeof b, "qe_bits_f64((qe_i64)0x%016llxULL)", (unsigned long long)bits);
    return b;
}

std::string and_valid(const std::string &a, const std::string &b)
{
    if (a == "true") return b;
    if (b == "true") return a;
    if (a == b) return a;
    return "(" + a + " && " + b + ")";
}

bool plain_literal(const SV &s)      // usable with the one-instruction comparison forms
{
    return s.is_const && s.k == s.k && s.k != 0.0;
}

const char *cmp_fn(int op)
{
    switch (op) {
    case OP_CMP_LT: return "qe_cmp_lt";
    case OP_CMP_LE: return "qe_cmp_le";
    case OP_CMP_GE: return "qe_cmp_ge";
    case OP_CMP_GT: return "qe_cmp_gt";
    case OP_CMP_EQ: return "qe_cmp_eq";
    default:        return "qe_cmp_ne";
    }
}

int cmp_swapped(int op)              // k OP x  ==  x OP' k
{
    switch (op) {
    case OP_CMP_LT: return OP_CMP_GT;
    case OP_CMP_LE: return OP_CMP_GE;
    case OP_CMP_GE: return OP_CMP_LE;
    case OP_CMP_GT: return OP_CMP_LT;
    default: return op;
    }
}

std::string slot_value(const BoundPipeline &bp, int slot)
{
    char b[48];
    const int t = bp.slot_info[slot].type;
    if (t == 3) snprintf(b, sizeof b, "qe_i64_to_f64(in.c%d)", slot);      // exact widening, |x| < 2^53
    else if (t == 4) snprintf(b, sizeof b, "qe_i32_to_f64(in.c%d)", slot);
    else snprintf(b, sizeof b, "in.c%d", slot);
    return b;
}

// Integer column against a literal: compare in the integer domain, no widening at all.
// For an integer x and a finite k:  x < k <=> x < ceil(k),  x <= k <=> x <= floor(k), and so on.
// (Literals that are NaN, -0.0 or out of the int64 range take the generic path.)
bool int_literal_ok(const SV &s)
{
    return s.is_const && s.k == s.k && std::fabs(s.k) < 4.0e18 && !(s.k == 0.0 && std::signbit(s.k));
}

// floor(k) / ceil(k) come from the kernel parameters (the host computes them per run, fill of QeParams::lit_lo /
// lit_hi) unless the literal had no parameter slot left; only `integral` -- part of the literal's class -- shapes the code.
std::string int_compare(int op, const std::string &x, const SV &lit)
{
    const double k = lit.k;
    const bool integral = std::floor(k) == std::ceil(k);
    char lo[48], hi[48], b[160];
    if (lit.lit >= 0) {
        snprintf(lo, sizeof lo, "p.lit_lo[%d]", lit.lit);
        snprintf(hi, sizeof hi, "p.lit_hi[%d]", lit.lit);
    } else {
        snprintf(lo, sizeof lo, "%lldLL", (long long)std::floor(k));
        snprintf(hi, sizeof hi, "%lldLL", (long long)std::ceil(k));
    }
    switch (op) {
    case OP_CMP_LT: snprintf(b, sizeof b, "((qe_i64)%s < %s)", x.c_str(), hi); break;
    case OP_CMP_LE: snprintf(b, sizeof b, "((qe_i64)%s <= %s)", x.c_str(), lo); break;
    case OP_CMP_GE: snprintf(b, sizeof b, "((qe_i64)%s >= %s)", x.c_str(), hi); break;
    case OP_CMP_GT: snprintf(b, sizeof b, "((qe_i64)%s > %s)", x.c_str(), lo); break;
    case OP_CMP_EQ:
        if (integral) snprintf(b, sizeof b, "((qe_i64)%s == %s)", x.c_str(), lo);
        else snprintf(b, sizeof b, "false");
        break;
    default:
        if (integral) snprintf(b, sizeof b, "((qe_i64)%s != %s)", x.c_str(), lo);
        else snprintf(b, sizeof b, "true");
        break;
    }
    return b;
}

// Emit straight-line code for one expression; returns its value / validity.
SV gen_expr(const BoundPipeline &bp, const BoundExpr &e, Gen &g)
{
    std::vector<SV> st;
    for (const Instr &in : e.code) {
        SV r;
        std::string key;
        if (in.op <= OP_CMP_NE) {                   // a function of the operands on top of the stack
            const int ar = op_arity(in.op);
            key = std::to_string((int)in.op);
            if (in.arg) key += "/" + std::to_string(in.arg);      // QE_CMP_IEEE: `a < b` in both orderings may appear in one pipeline
            for (size_t i = st.size() - (size_t)ar; i < st.size(); i++) key += "|" + st[i].val + "?" + st[i].valid;
            auto hit = g.memo.find(key);
            if (hit != g.memo.end()) {
                st.resize(st.size() - (size_t)ar);
                st.push_back(hit->second);
                continue;
            }
        }
        switch (in.op) {
        case OP_LOAD_COL: {
            const int slot = (int)in.arg;
            r.t = bp.slot_info[slot].type == 2 ? VT_BOOL : bp.slot_info[slot].type == 0 ? VT_STR : VT_F64;
            r.val = slot_value(bp, slot);              // STRING: the dictionary code itself (qe_i32)
            char b[32];
            if (bp.slot_info[slot].type == 3 || bp.slot_info[slot].type == 4) {
                snprintf(b, sizeof b, "in.c%d", slot);
                r.raw_int = b;
            }
            snprintf(b, sizeof b, "in.v%d", slot);
            r.valid = bp.slot_info[slot].nullable ? b : "true";
            break;
        }
        case OP_CONST_F64:                          // value from the kernel parameters: one kernel for every literal of a class
            r.t = VT_F64; r.valid = "true"; r.is_const = true; r.k = in.k; r.lit = in.lit;
            r.val = in.lit >= 0 ? "p.lit_f64[" + std::to_string(in.lit) + "]" : f64_literal(in.k);
            break;
        case OP_CONST_BOOL:
            r.t = VT_BOOL; r.val = in.arg ? "true" : "false"; r.valid = "true";
            break;
        case OP_CONST_STR:                          // StringLiteralExpression as its dictionary code
            r.t = VT_STR; r.valid = "true";
            r.val = in.lit >= 0 ? "((qe_i32)p.lit_lo[" + std::to_string(in.lit) + "])" : "((qe_i32)" + std::to_string(in.code) + ")";
            break;
        case OP_UNARY_PLUS: {                       // M/evaluator/Interpreter.kt:94
            r = st.back(); st.pop_back(); r.is_const = false; r.lit = -1; r.raw_int.clear();
            break;
        }
        case OP_UNARY_MINUS: {                      // Interpreter.kt:95
            SV a = st.back(); st.pop_back();
            r.t = VT_F64; r.valid = a.valid; r.val = g.tmp("t");
            g.line("const double " + r.val + " = qe_neg(" + a.val + ");");
            break;
        }
        case OP_NOT: {                              // ClosureCompiler.kt:115 / BytecodeCompiler.kt:346-351
            SV a = st.back(); st.pop_back();
            r.t = VT_BOOL; r.valid = a.valid; r.val = g.tmp("t");
            g.line("const bool " + r.val + " = !" + a.val + ";");
            break;
        }
        case OP_ADD: case OP_SUB: case OP_MUL: case OP_DIV: case OP_MOD: {   // Interpreter.kt:96-100
            SV b = st.back(); st.pop_back();
            SV a = st.back(); st.pop_back();
            const char *f = in.op == OP_ADD ? "qe_add" : in.op == OP_SUB ? "qe_sub" : in.op == OP_MUL ? "qe_mul"
                          : in.op == OP_DIV ? "qe_div" : "qe_mod";
            r.t = VT_F64; r.valid = and_valid(a.valid, b.valid); r.val = g.tmp("t");
            if (in.op == OP_MOD && b.is_const && !a.is_const) {
                // literal divisor: multiply by its reciprocal instead of dividing (qe_mod_k); both must be normal
                const double kk = std::fabs(b.k), inv = 1.0 / kk;
                if (kk >= 0x1p-1000 && kk <= 0x1p1000 && std::isfinite(inv)) {     // == Literal::mod_ok()
                    const std::string skk = b.lit >= 0 ? "p.lit_abs[" + std::to_string(b.lit) + "]" : f64_literal(kk);
                    const std::string sinv = b.lit >= 0 ? "p.lit_inv[" + std::to_string(b.lit) + "]" : f64_literal(inv);
                    g.line("const double " + r.val + " = qe_mod_k(" + a.val + ", " + b.val + ", " + skk + ", " + sinv + ");");
                    break;
                }
            }
            g.line("const double " + r.val + " = " + f + "(" + a.val + ", " + b.val + ");");
            break;
        }
        case OP_CMP_LT: case OP_CMP_LE: case OP_CMP_GE: case OP_CMP_GT: case OP_CMP_EQ: case OP_CMP_NE: {
            SV b = st.back(); st.pop_back();
            SV a = st.back(); st.pop_back();
            r.t = VT_BOOL; r.valid = and_valid(a.valid, b.valid); r.val = g.tmp("t");
            std::string rhs;
            const bool ieee = in.op <= OP_CMP_GT && (in.arg & 1u);      // QE_CMP_IEEE
            if (a.t == VT_BOOL || a.t == VT_STR) {  // Boolean / String equality (TypeCheck.kt:65-71); equal codes <=> equal strings
                rhs = std::string("(") + a.val + (in.op == OP_CMP_EQ ? " == " : " != ") + b.val + ")";
            } else if (ieee && !(int_literal_ok(b) && !a.raw_int.empty()) && !(int_literal_ok(a) && !b.raw_int.empty())) {
                // CLOSURE_COMPILER mode: (a as Double) < (b as Double) on unboxed doubles, ClosureCompiler.kt:127-130 --
                // every ordering comparison with a NaN is false, -0.0 equals 0.0.  (An integer column against a finite
                // literal other than -0.0 has neither case: the integer-domain forms below serve both orderings.)
                const char *o = in.op == OP_CMP_LT ? " < " : in.op == OP_CMP_LE ? " <= " : in.op == OP_CMP_GE ? " >= " : " > ";
                rhs = "(" + a.val + o + b.val + ")";
            } else if (int_literal_ok(b) && !a.raw_int.empty()) {
                rhs = int_compare(in.op, a.raw_int, b);
            } else if (int_literal_ok(a) && !b.raw_int.empty()) {
                rhs = int_compare(cmp_swapped(in.op), b.raw_int, a);
            } else if (plain_literal(b) && !a.is_const) {
                rhs = std::string(cmp_fn(in.op)) + "_k(" + a.val + ", " + b.val + ")";
            } else if (plain_literal(a) && !b.is_const) {
                rhs = std::string(cmp_fn(cmp_swapped(in.op))) + "_k(" + b.val + ", " + a.val + ")";
            } else {
                rhs = std::string(cmp_fn(in.op)) + "(" + a.val + ", " + b.val + ")";
            }
            g.line("const bool " + r.val + " = " + rhs + ";");
            break;
        }
        case OP_AND: case OP_OR: {                  // Kleene, Interpreter.kt:54-91
            SV b = st.back(); st.pop_back();
            SV a = st.back(); st.pop_back();
            r.t = VT_BOOL; r.val = g.tmp("t");
            if (a.valid == "true" && b.valid == "true") {
                r.valid = "true";
                g.line("const bool " + r.val + " = " + a.val + (in.op == OP_AND ? " && " : " || ") + b.val + ";");
            } else {
                r.valid = g.tmp("n");
                g.line("bool " + r.val + ", " + r.valid + ";");
                g.line(std::string(in.op == OP_AND ? "qe_and(" : "qe_or(") + a.val + ", " + a.valid + ", " + b.val + ", " +
                       b.valid + ", " + r.val + ", " + r.valid + ");");
            }
            break;
        }
        case OP_IF: {                               // Interpreter.kt:46-53
            SV f = st.back(); st.pop_back();
            SV t = st.back(); st.pop_back();
            SV c = st.back(); st.pop_back();
            r.t = t.t; r.val = g.tmp("t");
            g.line(std::string("const ") + (r.t == VT_F64 ? "double " : r.t == VT_STR ? "qe_i32 " : "bool ") + r.val + " = " + c.val + " ? " + t.val +
                   " : " + f.val + ";");
            std::string branch_valid = (t.valid == "true" && f.valid == "true")
                                           ? "true" : "(" + c.val + " ? " + t.valid + " : " + f.valid + ")";
            r.valid = and_valid(c.valid, branch_valid);
            if (r.valid != "true" && r.valid != c.valid) {
                std::string n = g.tmp("n");
                g.line("const bool " + n + " = " + r.valid + ";");
                r.valid = n;
            }
            break;
        }
        }
        if (!key.empty()) g.memo[key] = r;
        st.push_back(r);
    }
    return st.back();
}

const char *ctype_of(int qtype)
{
    switch (qtype) {
    case 1: return "double";
    case 2: return "bool";
    case 3: return "qe_i64";
    default: return "qe_i32";
    }
}

const char *suffix_of(int qtype) { return qtype == 1 ? "f64" : qtype == 3 ? "i64" : "i32"; }

// Layout of one shared-memory stage: per slot the values run, then (if nullable) the validity run.
struct StageLayout {
    std::vector<int> val_off, val_bytes, valid_off;
    int valid_bytes = 0, stride = 0, tx = 0;
};

StageLayout stage_layout(const BoundPipeline &bp, const TileShape &shape)
{
    StageLayout L;
    const int tile = shape.rows_per_tile();
    int off = 0;
    L.valid_bytes = tile / 8;
    for (const ColumnInfo &c : bp.slot_info) {
        const int w = c.type == 2 ? 0 : ((c.type == 4 || c.type == 0) ? 4 : 8);
        const int bytes = w ? tile * w : tile / 8;
        L.val_off.push_back(off);
        L.val_bytes.push_back(bytes);
        L.tx += bytes;
        off += (bytes + 127) / 128 * 128;
        if (c.nullable) {
            L.valid_off.push_back(off);
            L.tx += L.valid_bytes;
            off += (L.valid_bytes + 127) / 128 * 128;
        } else {
            L.valid_off.push_back(-1);
        }
    }
    L.stride = off ? off : 128;
    return L;
}

}  // namespace

size_t staged_smem_bytes(const BoundPipeline &bp, const TileShape &shape)
{
    if (!shape.staged && !shape.ring && !shape.greg && !shape.pring) return 0;
    return (size_t)stage_layout(bp, shape).stride * (size_t)shape.nstages;
}

size_t GroupPlan::copies_smem(int copies) const
{
    // qe_groupby_copies_body: [group][word][copy] accumulators + QeGroupCache2<NKEYS, CS, LMAX> (one 32-bit entry per
    // slot; key and table slot per group; four control words)
    const size_t nk = nkeys > 0 ? (size_t)nkeys : 1;
    const int ci = copies_index(copies);
    return (size_t)clmax[ci] * ((size_t)nwords * 8u * (size_t)copies + 8u * nk + 4u) + (size_t)ccslots[ci] * 4u + 16u;
}

size_t GroupPlan::shared_smemw() const
{
    const size_t nk = nkeys > 0 ? (size_t)nkeys : 1;
    // accumulators + QeGroupCache + one tag byte per warp and group
    return (size_t)slmaxw * (size_t)nwords * 8u * (size_t)(threads / 32) + (size_t)scslotsw * (8u * nk + 12u) + 4u * (size_t)slmaxw + 16u +
           (size_t)slmaxw * (size_t)(threads / 32);
}

void GroupPlan::scatter_shape(int rec_words, int &npart, int &pcap)
{
    // 512 partitions x 16 staging records for three-word records (196 KB); wider records: fewer records, then fewer partitions
    static const int cand[][2] = {{512, 16}, {512, 8}, {256, 8}, {128, 8}, {64, 8}, {32, 8}};
    for (const auto &c : cand) {
        npart = c[0];
        pcap = c[1];
        if (scatter_smem(rec_words, npart, pcap) <= (200u << 10)) return;
    }
}

GroupPlan plan_groupby(const BoundPipeline &bp, int threads, int lane_acc_kb)
{
    GroupPlan g;
    g.threads = threads;
    g.nkeys = bp.group_keys > 0 ? bp.group_keys : 0;
    auto word = [&](int kind, unsigned long long init) {
        g.kind.push_back(kind);
        g.init.push_back(init);
        return (int)g.kind.size() - 1;
    };
    word(5, 0xffffffff00000000ULL);          // first row (min, high half) | rows of the group (sum, low half)
    // Aggregates over the same input expression share words: SUM(x) and AVG(x) one sum, and every aggregate of
    // a nullable x one count (TPC-H Q1: sum(qty), avg(qty), sum(price), avg(price) -> two sum words, not four).
    auto expr_sig = [&](const BoundExpr &e) {
        std::string sg;
        char t[48];
        for (const Instr &in : e.code) {
            unsigned long long kb;
            memcpy(&kb, &in.k, sizeof kb);
            snprintf(t, sizeof t, "%u.%u.%llx.%d,", in.op, in.arg, in.op == OP_CONST_F64 ? kb : 0ULL, in.op == OP_CONST_STR ? in.code : 0);
            sg += t;
        }
        return sg;
    };
    std::map<std::string, int> shared;
    for (size_t k = 0; k < bp.agg_fns.size(); k++) {
        const int fn = bp.agg_fns[k];
        const BoundExpr &in = bp.proj[(size_t)g.nkeys + k];
        const std::string sg = expr_sig(in);
        auto shared_word = [&](const std::string &what, int kind, unsigned long long init) {
            auto it = shared.find(what + sg);
            if (it != shared.end()) return it->second;
            const int w = word(kind, init);
            shared[what + sg] = w;
            return w;
        };
        int vw = -1;
        if (fn == 2 || fn == 4) vw = shared_word("sum|", 0, 0);                                    // SUM / AVG: 0.0, Accumulators.kt:40,93
        else if (fn == 0) vw = shared_word("min|", 2, 0x7ff0000000000000ULL);                      // MIN: +Infinity, :57
        else if (fn == 1) vw = shared_word("max|", 3, 0xfff0000000000000ULL);                      // MAX: -Infinity, :75
        g.val_word.push_back(vw);
        g.cnt_word.push_back(in.nullable ? shared_word("cnt|", 1, 0) : -1);
    }
    g.nwords = (int)g.kind.size();
    {
        std::map<std::string, int> rec;
        for (size_t k = 0; k < bp.agg_fns.size(); k++) {
            const std::string sg = expr_sig(bp.proj[(size_t)g.nkeys + k]);
            auto it = rec.find(sg);
            if (it == rec.end()) it = rec.emplace(sg, (int)rec.size()).first;
            g.rec_word.push_back(1 + std::max(g.nkeys, 1) + it->second);
        }
        g.rec_words = 1 + std::max(g.nkeys, 1) + (int)rec.size();
    }
    const size_t per_group = (size_t)g.nwords * (size_t)threads * 8u;
    // shared memory of the private variant: LMAX groups of thread-private words + the key cache (>= 32 slots
    // per group, so that nearly every key sits in its first probe slot and the per-row path is straight-line
    // code; at most 1024 slots), together under 220 KB
    const int nk = g.nkeys > 0 ? g.nkeys : 1;
    for (g.lmax = (int)std::min<size_t>(64, (216u << 10) / per_group); g.lmax >= 1; g.lmax--) {
        g.cslots = 64;
        while (g.cslots < 32 * g.lmax && g.cslots < 1024) g.cslots *= 2;
        const size_t cache = (size_t)g.cslots * (8u * (size_t)nk + 12u) + 4u * (size_t)g.lmax + 64u;
        if ((size_t)g.lmax * per_group + cache <= (220u << 10)) break;
    }
    if (g.lmax < 1) { g.lmax = 0; g.cslots = 64; }
    // Shared-memory variants with 32 / 16 / 8 / 4 / 2 / 1 copies of the words per CTA (qe_groupby_copies_body): as many
    // groups as fit 224 KB with a key cache of >= 8 slots per group (a power of two; a slot is one 32-bit entry), then
    // 16 slots per group if they still fit: with buckets of two slots fewer than one key in a thousand sits outside its
    // home bucket, and only those rows (and new keys) leave the straight-line lookup.  (Round 1's cache took 20 bytes per
    // slot and four loads per lookup; with 8 slots per group a sixth of the keys sat past their first slot and the loop
    // those rows took made the kernel 6-18 x slower.)  `lane_acc_kb` caps the bytes of the copies (experiments; 0 = no cap).
    {
        const size_t budget = 224u << 10;
        const size_t acc_cap = (size_t)(lane_acc_kb > 0 ? lane_acc_kb : 224) << 10;
        for (int ci = 0; ci < 6; ci++) {
            const int copies = 1 << ci;
            int best = 0, best_cs = 64;
            for (int lm = 4096; lm >= 16; lm -= 8) {
                if ((size_t)lm * (size_t)g.nwords * 8u * (size_t)copies > acc_cap) continue;
                int cs = 64;
                while (cs < 8 * lm) cs *= 2;
                g.clmax[ci] = lm;
                g.ccslots[ci] = cs;
                if (g.copies_smem(copies) <= budget) { best = lm; best_cs = cs; break; }
            }
            if (copies == 1) {
                // the last stop before global atomics: 4 slots per group if that holds more groups (a few keys in a
                // hundred then sit outside their home bucket and are found by the in-line probe of the next slots)
                for (int lm = 4096; lm > best; lm -= 8) {
                    int cs = 64;
                    while (cs < 4 * lm) cs *= 2;
                    g.clmax[ci] = lm;
                    g.ccslots[ci] = cs;
                    if (g.copies_smem(copies) <= budget) { best = lm; best_cs = cs; break; }
                }
            }
            g.clmax[ci] = best;
            g.ccslots[ci] = best_cs;
            if (best > 0 && g.ccslots[ci] < 16 * best) {
                g.ccslots[ci] *= 2;
                if (g.copies_smem(copies) > budget) g.ccslots[ci] /= 2;
            }
        }
    }
    // one copy per warp (plain loads / stores, no atomics): threads / 32 copies, four key-cache slots per group while
    // they fit, two otherwise; up to 224 KB of the 227 KB a CTA may have
    // (>= 8 key-cache slots per group: with 4, a tenth of 1 024 keys sat past their home slot and every warp step took
    // the probing loop: 46 ms per 6e8 rows)
    for (g.slmaxw = 2048; g.slmaxw >= 64; g.slmaxw -= 32) {
        g.scslotsw = 64;
        while (g.scslotsw < 8 * g.slmaxw) g.scslotsw *= 2;
        if (g.shared_smemw() <= (224u << 10)) break;
    }
    if (g.slmaxw < 64) { g.slmaxw = 0; g.scslotsw = 64; }
    // register variant: per group two 32-bit registers (first row, rows) + two per further word; ~96 registers
    // of accumulators leave room for the row and the expression temporaries under the 168 a 384-thread CTA (1 producer + 11 consumer warps) gets
    const int reg_per_group = 2 + 2 * (g.nwords - 1);
    g.rmax = std::min(16, 96 / reg_per_group);
    g.rmax_regs = g.rmax < 2 ? 0 : g.rmax;
    // ... and it pays only where the row is wide enough for the issue slots: a row costs about
    // 100 + RMAX * (4 + words) instructions (every group's words are touched, predicated), and at the HBM rate an
    // SM has ~5.7 issue slots per row byte (measured: Q1 shape, 48-byte rows, 6 words, RMAX 8: 5.8 ms against 7.3
    // for the private variant; SUM, COUNT over 16-byte rows: 8.1 ms against 2.0)
    int row_bytes = 0;
    for (const ColumnInfo &c : bp.slot_info) row_bytes += c.type == 2 ? 1 : ((c.type == 4 || c.type == 0) ? 4 : 8);
    const int fit = (5 * row_bytes - 100) / (4 + g.nwords);
    g.rmax = std::min(g.rmax, fit);
    if (g.rmax < 4) g.rmax = 0;
    return g;
}

namespace {

// An expression as a tree again (the bound code is postfix), for the two questions below.
struct ExprTree {
    struct Node { int op; uint32_t arg; int kid[3]; };
    std::vector<Node> nodes;
    int root = -1;
    explicit ExprTree(const BoundExpr &e)
    {
        std::vector<int> st;
        for (const Instr &in : e.code) {
            Node n{in.op, in.arg, {-1, -1, -1}};
            const int ar = op_arity(in.op);
            for (int i = ar - 1; i >= 0; i--) { n.kid[i] = st.back(); st.pop_back(); }
            nodes.push_back(n);
            st.push_back((int)nodes.size() - 1);
        }
        root = st.empty() ? -1 : st.back();
    }
};

// Columns (slots) that are certainly non-null in every row the filter keeps.  FilterOperator keeps a row iff the
// predicate is non-null and true (M/operator/FilterOperator.kt:20).  A Kleene AND is true only if both sides are true
// (Interpreter.kt:54-72); every function but AND / OR / IF is null as soon as one operand is (Interpreter.kt:35-42), so a
// non-null result means non-null operands; IF needs a non-null condition (Interpreter.kt:46-53).  OR promises nothing.
void slots_nonnull(const ExprTree &t, int n, std::vector<bool> &nn)
{
    const ExprTree::Node &x = t.nodes[(size_t)n];
    if (x.op == OP_LOAD_COL) { nn[x.arg] = true; return; }
    if (x.op == OP_AND || x.op == OP_OR || op_arity(x.op) == 0) return;
    if (x.op == OP_IF) { slots_nonnull(t, x.kid[0], nn); return; }
    for (int i = 0; i < op_arity(x.op); i++) slots_nonnull(t, x.kid[i], nn);
}
void slots_of_true(const ExprTree &t, int n, std::vector<bool> &nn)
{
    const ExprTree::Node &x = t.nodes[(size_t)n];
    if (x.op == OP_AND) { slots_of_true(t, x.kid[0], nn); slots_of_true(t, x.kid[1], nn); return; }
    slots_nonnull(t, n, nn);
}

// Is the expression null exactly when one of the columns it reads is (no AND / OR / IF), and which columns are those?
bool strict_slots(const BoundExpr &e, std::vector<bool> &used)
{
    for (const Instr &in : e.code) {
        if (in.op == OP_AND || in.op == OP_OR || in.op == OP_IF) return false;
        if (in.op == OP_LOAD_COL) used[in.arg] = true;
    }
    return true;
}

}  // namespace

// Outputs whose validity the filter already decides: every kept row has them non-null, so their compacted validity
// bitmap is a run of ones as long as the output (written per tile by the ring kernel's look-back warp instead of bit by
// bit by the warps that scatter).  `x < k` over a nullable x returning x is the plain case.
std::vector<bool> implied_valid_outputs(const BoundPipeline &bp)
{
    std::vector<bool> implied(bp.proj.size(), false);
    if (!bp.has_filter) return implied;
    std::vector<bool> nn(bp.slot_info.size(), false);
    const ExprTree ft(bp.filter);
    if (ft.root < 0) return implied;
    slots_of_true(ft, ft.root, nn);
    for (size_t k = 0; k < bp.proj.size(); k++) {
        if (!bp.proj[k].nullable) continue;
        std::vector<bool> used(bp.slot_info.size(), false);
        if (!strict_slots(bp.proj[k], used)) continue;
        bool ok = true;
        for (size_t sl = 0; sl < used.size(); sl++)
            if (used[sl] && bp.slot_info[sl].nullable && !nn[sl]) ok = false;
        implied[k] = ok;
    }
    return implied;
}

std::string generate_source(const BoundPipeline &bp, KernelMode mode, const TileShape &shape)
{
    std::string s;
    char b[256];
    const int nslots = (int)bp.slot_info.size();
    const int nproj = (int)bp.proj.size();
    // validity bitmaps the ring filter writes as runs of ones (see implied_valid_outputs); the other kernels write every bit
    std::vector<bool> implied(bp.proj.size(), false);
    if (mode == KM_FILTER && shape.ring) implied = implied_valid_outputs(bp);

    s += "// generated by qe_codegen.cpp -- pipeline: " + bp.signature() + "\n";
    s += "#include \"qe_device.cuh\"\n\nstruct P {\n";
    // bit columns of the outputs (validity bitmaps and BOOLEAN values)
    struct BitCol { std::string ptr, bit; };
    std::vector<BitCol> bitcols;
    const bool is_group = mode == KM_GROUPBY_PRIVATE || mode == KM_GROUPBY_ATOMIC || mode == KM_GROUPBY_SHARED || mode == KM_GROUPBY_REG;
    if (mode != KM_AGG && !is_group) {
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].has_validity && !implied[(size_t)k]) {
                snprintf(b, sizeof b, "p.out_valid[%d]", k);
                BitCol bc; bc.ptr = b;
                snprintf(b, sizeof b, "o.n%d", k); bc.bit = b;
                bitcols.push_back(bc);
            }
            if (bp.outs[k].type == 2) {
                snprintf(b, sizeof b, "(qe_u32 *)p.out_values[%d]", k);
                BitCol bc; bc.ptr = b;
                if (bp.outs[k].has_validity) snprintf(b, sizeof b, "(o.o%d && o.n%d)", k, k);
                else snprintf(b, sizeof b, "o.o%d", k);
                bc.bit = b;
                bitcols.push_back(bc);
            }
        }
    }
    snprintf(b, sizeof b, "    static constexpr int STEPS = %d, VEC = %d, NBITCOLS = %d;\n", shape.steps, shape.vec,
             (int)bitcols.size());
    s += b;

    // ---- Row
    s += "    struct Row {\n";
    for (int i = 0; i < nslots; i++) {
        snprintf(b, sizeof b, "        %s c%d;\n", ctype_of(bp.slot_info[i].type), i); s += b;
        if (bp.slot_info[i].nullable) { snprintf(b, sizeof b, "        bool v%d;\n", i); s += b; }
    }
    if (nslots == 0) s += "        int unused;\n";
    s += "    };\n";
    // ---- Out
    s += "    struct Out {\n";
    for (int k = 0; k < nproj; k++) {
        snprintf(b, sizeof b, "        %s o%d;\n", ctype_of(bp.outs[k].type), k); s += b;
        if (bp.proj[k].nullable) { snprintf(b, sizeof b, "        bool n%d;\n", k); s += b; }
    }
    if (nproj == 0) s += "        int unused;\n";
    s += "    };\n";

    // ---- load: one vector load per column, MemorySourceOperator.next (M/operator/MemorySourceOperator.kt:18-36)
    s += "    static __device__ __forceinline__ void load(const QeParams &p, qe_i64 r, Row (&rows)[VEC]) {\n";
    for (int i = 0; i < nslots; i++) {
        const int t = bp.slot_info[i].type;
        if (t == 2) {
            snprintf(b, sizeof b,
                     "        { const qe_u32 m = qe_load_bits<VEC>((const qe_u32 *)p.col_values[%d], r);\n"
                     "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].c%d = (m >> j) & 1u; }\n", i, i);
        } else {
            snprintf(b, sizeof b,
                     "        { %s t[VEC]; qe_load_%s<VEC>(p.col_values[%d], r, t);\n"
                     "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].c%d = t[j]; }\n",
                     ctype_of(t), suffix_of(t), i, i);
        }
        s += b;
        if (bp.slot_info[i].nullable) {
            snprintf(b, sizeof b,
                     "        { const qe_u32 m = qe_load_bits<VEC>(p.col_valid[%d], r);\n"
                     "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].v%d = (m >> j) & 1u; }\n", i, i);
            s += b;
        }
    }
    s += "    }\n";

    if ((mode == KM_FILTER && (shape.staged || shape.ring)) || (mode == KM_GROUPBY_REG && shape.greg) || (mode == KM_PROJECT && shape.pring)) {
        const bool ringlike = shape.ring || shape.greg || shape.pring;
        const StageLayout L = stage_layout(bp, shape);
        snprintf(b, sizeof b, "    static constexpr int NSTAGES = %d, STAGE_BYTES = %d, STAGE_TX = %d;\n", shape.nstages, L.stride, L.tx);
        s += b;
        if (ringlike) {
            snprintf(b, sizeof b, "    static constexpr int TILE_ROWS = %d, NCW = %d, NLB = %d, DEFER = %d, LOADS = %d, LIST_CAP = %d, NSW = %d;\n", shape.tile_rows, shape.ncw,
                     shape.nlb, shape.defer, shape.loads, shape.ring ? shape.list_cap : 0, (shape.ring && shape.list_cap) ? shape.nsw : 0);
            s += b;
            // one row (lane-strided: a warp step reads 32 consecutive rows, conflict-free) from a stage
            s += "    static __device__ __forceinline__ void load_row(qe_u32 stage, int wrow0, int c, Row &row) {\n";
            for (int i = 0; i < nslots; i++) {
                const int t = bp.slot_info[i].type;
                if (t == 2) snprintf(b, sizeof b, "        row.c%d = qe_stage1_bit(stage + %d, wrow0, c);\n", i, L.val_off[i]);
                else snprintf(b, sizeof b, "        row.c%d = qe_stage1_%s(stage + %d, wrow0, c);\n", i, suffix_of(t), L.val_off[i]);
                s += b;
                if (bp.slot_info[i].nullable) {
                    snprintf(b, sizeof b, "        row.v%d = qe_stage1_bit(stage + %d, wrow0, c);\n", i, L.valid_off[i]);
                    s += b;
                }
            }
            s += "    }\n";
            // row r of the tile, any r (the sparse path of the ring kernel: one kept row per lane)
            s += "    static __device__ __forceinline__ void load_at(qe_u32 stage, int r, Row &row) {\n";
            for (int i = 0; i < nslots; i++) {
                const int t = bp.slot_info[i].type;
                if (t == 2) snprintf(b, sizeof b, "        row.c%d = qe_stage_at_bit(stage + %d, r);\n", i, L.val_off[i]);
                else snprintf(b, sizeof b, "        row.c%d = qe_stage_at_%s(stage + %d, r);\n", i, suffix_of(t), L.val_off[i]);
                s += b;
                if (bp.slot_info[i].nullable) {
                    snprintf(b, sizeof b, "        row.v%d = qe_stage_at_bit(stage + %d, r);\n", i, L.valid_off[i]);
                    s += b;
                }
            }
            s += "    }\n";
        }
        // one thread: a bulk copy per column run of the tile, all signalling the stage's mbarrier
        const char *bulk = ringlike ? "qe_bulk_g2s_stream" : "qe_bulk_g2s";
        const char *pol = ringlike ? ", policy" : "";
        if (ringlike)
            s += "    static __device__ __forceinline__ void issue(const QeParams &p, qe_i64 row0, qe_u32 stage, qe_u32 mbar, qe_u64 policy) {\n";
        else
        s += "    static __device__ __forceinline__ void issue(const QeParams &p, qe_i64 row0, qe_u32 stage, qe_u32 mbar) {\n";
        for (int i = 0; i < nslots; i++) {
            const int t = bp.slot_info[i].type;
            if (t == 2)
                snprintf(b, sizeof b, "        %s(stage + %d, (const char *)p.col_values[%d] + (row0 >> 3), %d, mbar%s);\n",
                         bulk, L.val_off[i], i, L.val_bytes[i], pol);
            else
                snprintf(b, sizeof b, "        %s(stage + %d, (const char *)p.col_values[%d] + row0 * %d, %d, mbar%s);\n",
                         bulk, L.val_off[i], i, (t == 4 || t == 0) ? 4 : 8, L.val_bytes[i], pol);
            s += b;
            if (bp.slot_info[i].nullable) {
                snprintf(b, sizeof b, "        %s(stage + %d, (const char *)p.col_valid[%d] + (row0 >> 3), %d, mbar%s);\n",
                         bulk, L.valid_off[i], i, L.valid_bytes, pol);
                s += b;
            }
        }
        s += "    }\n";
        if (shape.staged) {
        s += "    static __device__ __forceinline__ void load_stage(const unsigned char *stage, int r, Row (&rows)[VEC]) {\n";
        for (int i = 0; i < nslots; i++) {
            const int t = bp.slot_info[i].type;
            if (t == 2) {
                snprintf(b, sizeof b,
                         "        { const qe_u32 m = qe_stage_bits(stage + %d, r);\n"
                         "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].c%d = (m >> j) & 1u; }\n", L.val_off[i], i);
            } else {
                snprintf(b, sizeof b,
                         "        { %s t[VEC]; qe_stage_%s(stage + %d, r, t);\n"
                         "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].c%d = t[j]; }\n",
                         ctype_of(t), suffix_of(t), L.val_off[i], i);
            }
            s += b;
            if (bp.slot_info[i].nullable) {
                snprintf(b, sizeof b,
                         "        { const qe_u32 m = qe_stage_bits(stage + %d, r);\n"
                         "          _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) rows[j].v%d = (m >> j) & 1u; }\n", L.valid_off[i], i);
                s += b;
            }
        }
        s += "    }\n";
        }
    }

    // ---- eval: RowCallable.invoke for the filter and every projection
    s += "    static __device__ __forceinline__ bool eval(const QeParams &p, const Row &in, Out &o) {\n";
    {
        Gen g;
        std::string keep = "true";
        if (bp.has_filter) {
            SV f = gen_expr(bp, bp.filter, g);
            keep = and_valid(f.val, f.valid);        // FilterOperator.kt:20: non-null AND true
        }
        for (int k = 0; k < nproj; k++) {
            const BoundExpr &e = bp.proj[k];
            if (e.identity_slot >= 0) {
                snprintf(b, sizeof b, "o.o%d = in.c%d;", k, e.identity_slot); g.line(b);
                if (e.nullable) { snprintf(b, sizeof b, "o.n%d = in.v%d;", k, e.identity_slot); g.line(b); }
            } else {
                SV v = gen_expr(bp, e, g);
                snprintf(b, sizeof b, "o.o%d = ", k); g.line(std::string(b) + v.val + ";");
                if (e.nullable) { snprintf(b, sizeof b, "o.n%d = ", k); g.line(std::string(b) + v.valid + ";"); }
            }
        }
        s += g.body;
        s += "        return " + keep + ";\n";
    }
    s += "    }\n";

    if (mode == KM_FILTER || (mode == KM_PROJECT && shape.pring)) {
        s += "    static __device__ __forceinline__ void emit(const QeParams &p, qe_u64 pos, const Out &o) {\n";
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].type == 2) continue;
            snprintf(b, sizeof b, "        ((%s *)p.out_values[%d])[pos] = o.o%d;\n", ctype_of(bp.outs[k].type), k, k);
            s += b;
        }
        s += "    }\n";
    }
    if (mode == KM_FILTER && shape.ring) {
        // per-warp output pointers (advanced to the warp's first output once per tile) + 32-bit indices
        s += "    struct Ptrs {";
        bool any = false;
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].type == 2) continue;
            snprintf(b, sizeof b, " %s *o%d;", ctype_of(bp.outs[k].type), k); s += b;
            any = true;
        }
        if (!any) s += " int unused;";
        s += " };\n";
        s += "    static __device__ __forceinline__ Ptrs out_ptrs(const QeParams &p, qe_u64 base) {\n        Ptrs q;\n";
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].type == 2) continue;
            snprintf(b, sizeof b, "        q.o%d = qe_opaque_ptr((%s *)p.out_values[%d] + base);\n", k, ctype_of(bp.outs[k].type), k); s += b;
        }
        if (!any) s += "        q.unused = 0;\n";
        s += "        return q;\n    }\n";
        s += "    static __device__ __forceinline__ void emit_ptr(const Ptrs &q, qe_u32 idx, const Out &o, bool ok) {\n";
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].type == 2) continue;
            snprintf(b, sizeof b, "        qe_stg_if(ok, q.o%d + idx, o.o%d);\n", k, k); s += b;
        }
        s += "    }\n";
    }
    if (mode == KM_PROJECT) {
        s += "    static __device__ __forceinline__ void emit_dense(const QeParams &p, qe_i64 r, const Out (&o)[VEC], qe_u32 inb) {\n";
        for (int k = 0; k < nproj; k++) {
            if (bp.outs[k].type == 2) continue;
            snprintf(b, sizeof b,
                     "        { %s t[VEC]; _Pragma(\"unroll\") for (int j = 0; j < VEC; j++) t[j] = o[j].o%d;\n"
                     "          qe_store_%s<VEC>(p.out_values[%d], r, t); }\n",
                     ctype_of(bp.outs[k].type), k, suffix_of(bp.outs[k].type), k);
            s += b;
        }
        s += "    }\n";
    }
    if (mode == KM_FILTER && shape.ring) {
        int n = 0;
        std::string body;
        for (int k = 0; k < nproj; k++) {
            if (!(bp.outs[k].has_validity && implied[(size_t)k])) continue;
            snprintf(b, sizeof b, "        case %d: return p.out_valid[%d];\n", n++, k);
            body += b;
        }
        snprintf(b, sizeof b, "    static constexpr int NIMPLIED = %d;      // validity bitmaps that are all ones: the filter keeps no row where they are null\n", n);
        s += b;
        s += "    static __device__ __forceinline__ qe_u32 *implied_col(const QeParams &p, int i) {\n        switch (i) {\n" + body +
             "        }\n        return 0;\n    }\n";
    }
    if (mode != KM_AGG && !is_group) {
        s += "    static __device__ __forceinline__ qe_u32 *bitcol(const QeParams &p, int b) {\n        switch (b) {\n";
        for (size_t i = 0; i < bitcols.size(); i++) {
            snprintf(b, sizeof b, "        case %d: return ", (int)i); s += b; s += bitcols[i].ptr + ";\n";
        }
        s += "        }\n        return 0;\n    }\n";
        s += "    static __device__ __forceinline__ bool bit(const Out &o, int b) {\n        switch (b) {\n";
        for (size_t i = 0; i < bitcols.size(); i++) {
            snprintf(b, sizeof b, "        case %d: return ", (int)i); s += b; s += bitcols[i].bit + ";\n";
        }
        s += "        }\n        return false;\n    }\n";
    }
    if (mode == KM_AGG) {
        // GlobalAggregationOperator.open (M/operator/GlobalAggregationOperator.kt:10-25): accumulator k takes
        // projection k, nulls are skipped
        s += "    struct Accs {";
        for (size_t k = 0; k < bp.agg_fns.size(); k++) { snprintf(b, sizeof b, " QeAcc a%d;", (int)k); s += b; }
        s += " };\n";
        s += "    static __device__ __forceinline__ void acc_init(Accs &a) {";
        for (size_t k = 0; k < bp.agg_fns.size(); k++) { snprintf(b, sizeof b, " qe_acc_init(a.a%d);", (int)k); s += b; }
        s += " }\n";
        s += "    static __device__ __forceinline__ void acc_add(const QeParams &p, Accs &a, const Out &o, bool keep) {\n";
        for (size_t k = 0; k < bp.agg_fns.size(); k++) {
            std::string val = "0.0";
            if (bp.outs[k].type == 1) { snprintf(b, sizeof b, "o.o%d", (int)k); val = b; }
            else if (bp.outs[k].type == 3) { snprintf(b, sizeof b, "qe_i64_to_f64(o.o%d)", (int)k); val = b; }
            else if (bp.outs[k].type == 4) { snprintf(b, sizeof b, "qe_i32_to_f64(o.o%d)", (int)k); val = b; }
            std::string on = "keep";
            if (bp.proj[k].nullable) { snprintf(b, sizeof b, "(keep && o.n%d)", (int)k); on = b; }
            snprintf(b, sizeof b, "        qe_acc_add<%d>(a.a%d, ", bp.agg_fns[k], (int)k);
            s += b; s += val + ", " + on + ");\n";
        }
        s += "    }\n";
        s += "    static __device__ __forceinline__ void acc_store(const QeParams &p, Accs &a, QeAcc *slot, qe_i64 part) {\n";
        for (size_t k = 0; k < bp.agg_fns.size(); k++) {
            snprintf(b, sizeof b, "        qe_acc_block_reduce_store<%d>(a.a%d, slot, p.agg_partials, part, %d, %d);\n", bp.agg_fns[k],
                     (int)k, (int)bp.agg_fns.size(), (int)k);
            s += b;
        }
        s += "    }\n";
    }
    if (is_group) {
        // GroupByAggregationOperator.open (M/operator/GroupByAggregationOperator.kt:21-48): key = the first
        // NKEYS values of the row, accumulator i takes value NKEYS + i, nulls are skipped
        GroupPlan gp = plan_groupby(bp, shape.gthreads, shape.lane_acc_kb);
        if (shape.rmax_hint > 0) gp.rmax = std::min(shape.rmax_hint, gp.rmax_regs);
        snprintf(b, sizeof b, "    static constexpr int NKEYS = %d, NWORDS = %d, LMAX = %d, CSLOTS = %d, GTHREADS = %d;\n", gp.nkeys > 0 ? gp.nkeys : 1,
                 gp.nwords, gp.lmax > 0 ? gp.lmax : 1, gp.cslots, gp.threads);
        s += b;
        if (shape.scopies < 0)          // one copy per warp, plain loads and stores (qe_groupby_shared_body, WARPC)
            snprintf(b, sizeof b, "    static constexpr int SLMAX = %d, SCSLOTS = %d, SCOPIES = 0;\n", gp.slmaxw > 0 ? gp.slmaxw : 64, gp.scslotsw);
        else {                          // 32 / 16 / 8 / 4 / 2 / 1 copies per CTA (qe_groupby_copies_body)
            const int ci = GroupPlan::copies_index(shape.scopies);
            snprintf(b, sizeof b, "    static constexpr int SLMAX = %d, SCSLOTS = %d, SCOPIES = %d;\n", gp.clmax[ci] > 0 ? gp.clmax[ci] : 16,
                     gp.clmax[ci] > 0 ? gp.ccslots[ci] : 64, 1 << ci);
        }
        s += b;
        {
            int rcs = 64;
            while (rcs < 32 * gp.rmax) rcs *= 2;
            snprintf(b, sizeof b, "    static constexpr int RMAX = %d, RCSLOTS = %d;\n", gp.rmax > 0 ? gp.rmax : 1, rcs);
            s += b;
        }
        s += "    static __device__ __forceinline__ void key_pack(const Out &o, qe_u64 (&k)[NKEYS], qe_u32 &nulls) {\n        nulls = 0u;\n";
        if (gp.nkeys == 0) s += "        k[0] = 0ull;\n";
        for (int i = 0; i < gp.nkeys; i++) {
            const int t = bp.outs[i].type;
            if (t == 1) snprintf(b, sizeof b, "        k[%d] = qe_key_f64(o.o%d);\n", i, i);
            else if (t == 2) snprintf(b, sizeof b, "        k[%d] = o.o%d ? 1ull : 0ull;\n", i, i);
            else snprintf(b, sizeof b, "        k[%d] = (qe_u64)(qe_i64)o.o%d;\n", i, i);
            s += b;
            if (bp.proj[i].nullable) {
                snprintf(b, sizeof b, "        if (!o.n%d) { k[%d] = 0ull; nulls |= %uu; }\n", i, i, 1u << i);
                s += b;
            }
        }
        s += "    }\n";
        s += "    static __device__ __forceinline__ qe_u64 word_init(int w) {\n        switch (w) {\n";
        for (int w = 0; w < gp.nwords; w++) {
            snprintf(b, sizeof b, "        case %d: return 0x%llxull;\n", w, gp.init[w]); s += b;
        }
        s += "        }\n        return 0ull;\n    }\n";
        s += "    static __device__ __forceinline__ int word_kind(int w) {\n        switch (w) {\n";
        for (int w = 0; w < gp.nwords; w++) {
            snprintf(b, sizeof b, "        case %d: return %d;\n", w, gp.kind[w]); s += b;
        }
        s += "        }\n        return 1;\n    }\n";
        // value and validity of aggregate input j
        auto agg_val = [&](int j) {
            const int k = gp.nkeys + j;
            char v[64];
            if (bp.outs[k].type == 1) snprintf(v, sizeof v, "o.o%d", k);
            else if (bp.outs[k].type == 3) snprintf(v, sizeof v, "qe_i64_to_f64(o.o%d)", k);
            else if (bp.outs[k].type == 4) snprintf(v, sizeof v, "qe_i32_to_f64(o.o%d)", k);
            else snprintf(v, sizeof v, "0.0");
            return std::string(v);
        };
        for (int variant = 0; variant < 4; variant++) {
            // 0 private (shared-memory words), 1 global atomics, 2 registers with selects, 3 registers with DFMA by a
            // 0.0 / 1.0 mask (finite inputs only: inf * 0.0 would poison the other groups; all_finite() guards it)
            const bool atomic = variant == 1, regs = variant >= 2, rfma = variant == 3;
            // register variant: called for every g in [0, RMAX), branch-free -- every update is a predicated
            // instruction or a select on (l == g).  (As `if (l == g) update` the compiler built a divergent switch:
            // every group met by the warp ran as its own branch, ~140 cycles per group and warp step.)
            s += atomic ? "    template <int AS, bool FIRST = true> static __device__ __forceinline__ void acc_atomic(qe_u64 *a, const Out &o, qe_u64 row) {\n"
               : rfma ? "    static __device__ __forceinline__ void acc_regs_fma(qe_u32 &first, qe_u32 &rows, qe_u64 (&a)[NWORDS], const Out &o, qe_u32 row, int l, int g) {\n"
               : regs ? "    static __device__ __forceinline__ void acc_regs(qe_u32 &first, qe_u32 &rows, qe_u64 (&a)[NWORDS], const Out &o, qe_u32 row, int l, int g) {\n"
                        : "    template <int WS> static __device__ __forceinline__ void acc_plain(qe_u64 *a, const Out &o, qe_u64 row) {\n";
            auto at = [&](int w) {
                char a[48];
                if (atomic) snprintf(a, sizeof a, "a + %d * AS", w);
                else if (regs) snprintf(a, sizeof a, "a[%d]", w);
                else snprintf(a, sizeof a, "a[%d * WS]", w);
                return std::string(a);
            };
            if (atomic) {
                // word 0, little endian: rows in the low half, first row in the high half.  After the first few rows
                // of a group nothing is smaller any more -- look before the atomic.
                // (FIRST = false: the caller knows whether the group is young enough for this row to matter, qe_groupby_copies_body)
                s += "        if (FIRST && (qe_u32)row < ((volatile qe_u32 *)a)[1]) atomicMin(((unsigned int *)a) + 1, (unsigned int)row);\n";
                s += "        atomicAdd((unsigned int *)a, 1u);\n";
            } else if (regs) {
                // a thread meets its rows in ascending order, but min() keeps this independent of that
                s += "        const bool m = l == g;\n"
                     "        first = min(first, m ? row : 0xffffffffu);\n        rows += m ? 1u : 0u;\n";
                if (rfma) s += "        const double mg = m ? 1.0 : 0.0;\n";
            } else {
                s += "        { const qe_u64 w0 = a[0]; const qe_u64 f = w0 >> 32;\n"
                     "          a[0] = ((row < f ? row : f) << 32) | ((w0 + 1ull) & 0xffffffffull); }\n";
            }
            std::vector<bool> done((size_t)gp.nwords, false);   // a shared word is updated by the first aggregate that uses it
            for (size_t j = 0; j < bp.agg_fns.size(); j++) {
                const int fn = bp.agg_fns[j];
                const int k = gp.nkeys + (int)j;
                const int vw = gp.val_word[j], cw = gp.cnt_word[j];
                std::string guard;
                if (bp.proj[k].nullable) { snprintf(b, sizeof b, "if (o.n%d) ", k); guard = b; }
                std::string stmts;
                const std::string v = agg_val((int)j);
                if (regs) {
                    // l of this aggregate: -1 when its input is null (nulls are skipped, Accumulators.kt)
                    std::string lj = "l";
                    if (bp.proj[k].nullable) { snprintf(b, sizeof b, "(o.n%d ? l : -1)", k); lj = b; }
                    if (vw >= 0 && !done[(size_t)vw]) {
                        done[(size_t)vw] = true;
                        if ((fn == 2 || fn == 4) && rfma) {
                            std::string mj = "mg";
                            if (bp.proj[k].nullable) { snprintf(b, sizeof b, "((o.n%d && m) ? 1.0 : 0.0)", k); mj = b; }
                            stmts += at(vw) + " = (qe_u64)qe_f64_bits(fma(" + v + ", " + mj + ", qe_bits_f64((qe_i64)" + at(vw) + "))); ";
                        } else {
                            const char *f = (fn == 2 || fn == 4) ? "qe_add" : fn == 0 ? "qe_math_min" : "qe_math_max";
                            stmts += at(vw) + " = (" + lj + " == g) ? (qe_u64)qe_f64_bits(" + f +
                                     "(qe_bits_f64((qe_i64)" + at(vw) + "), " + v + ")) : " + at(vw) + "; ";
                        }
                    }
                    if (cw >= 0 && !done[(size_t)cw]) {
                        done[(size_t)cw] = true;
                        stmts += at(cw) + " += (" + lj + " == g) ? 1ull : 0ull; ";
                    }
                    if (!stmts.empty()) s += "        " + stmts + "\n";
                    continue;
                }
                if (vw >= 0 && !done[(size_t)vw]) {
                    done[(size_t)vw] = true;
                    if (atomic) {
                        if (fn == 2 || fn == 4) stmts += "atomicAdd((double *)(" + at(vw) + "), " + v + "); ";
                        else if (fn == 0) stmts += "qe_atomic_min_f64(" + at(vw) + ", " + v + "); ";
                        else stmts += "qe_atomic_max_f64(" + at(vw) + ", " + v + "); ";
                    } else {
                        const char *f = (fn == 2 || fn == 4) ? "qe_add" : fn == 0 ? "qe_math_min" : "qe_math_max";
                        stmts += at(vw) + " = (qe_u64)qe_f64_bits(" + f + "(qe_bits_f64((qe_i64)" + at(vw) + "), " + v + ")); ";
                    }
                }
                if (cw >= 0 && !done[(size_t)cw]) {
                    done[(size_t)cw] = true;
                    // a table has < 2^32 rows: the count fits the low half, and 32-bit adds are native in shared memory too
                    if (atomic) stmts += "atomicAdd((unsigned int *)(" + at(cw) + "), 1u); ";
                    else stmts += at(cw) + " += 1ull; ";
                }
                if (!stmts.empty()) s += "        " + guard + "{ " + stmts + "}\n";
            }
            s += "    }\n";
            if (variant == 0)      // thread-private copies: word stride = groups x threads
                s += "    static __device__ __forceinline__ void acc_private(qe_u64 *a, const Out &o, qe_u64 row) { acc_plain<LMAX * GTHREADS>(a, o, row); }\n";
        }
        // two-pass variant: the record of a row (see GroupPlan::rec_words)
        {
            snprintf(b, sizeof b, "    static constexpr int RW = %d, NPART = %d, PCAP = %d;\n", gp.rec_words, shape.npart > 0 ? shape.npart : 32,
                     shape.pcap > 0 ? shape.pcap : 8);
            s += b;
            auto to_word = [&](int k) {
                char v[96];
                const int t = bp.outs[k].type;
                if (t == 1) snprintf(v, sizeof v, "(qe_u64)qe_f64_bits(o.o%d)", k);
                else if (t == 2) snprintf(v, sizeof v, "(o.o%d ? 1ull : 0ull)", k);
                else snprintf(v, sizeof v, "(qe_u64)(qe_i64)o.o%d", k);
                return std::string(v);
            };
            auto from_word = [&](int k, int w) {
                char v[96];
                const int t = bp.outs[k].type;
                if (t == 1) snprintf(v, sizeof v, "qe_bits_f64((qe_i64)rec[%d])", w);
                else if (t == 2) snprintf(v, sizeof v, "rec[%d] != 0ull", w);
                else if (t == 3) snprintf(v, sizeof v, "(qe_i64)rec[%d]", w);
                else snprintf(v, sizeof v, "(qe_i32)(qe_i64)rec[%d]", w);
                return std::string(v);
            };
            const int nk = gp.nkeys > 0 ? gp.nkeys : 1;
            s += "    static __device__ __forceinline__ void record_pack(const Out &o, qe_u32 row, const qe_u64 (&k)[NKEYS], qe_u32 nulls, qe_u64 (&rec)[RW]) {\n"
                 "        qe_u64 m = (qe_u64)row | ((qe_u64)nulls << 32);\n";
            for (int i = 0; i < nk; i++) { snprintf(b, sizeof b, "        rec[%d] = k[%d];\n", 1 + i, i); s += b; }
            std::vector<bool> packed((size_t)gp.rec_words, false);
            for (size_t j = 0; j < bp.agg_fns.size(); j++) {
                const int k = gp.nkeys + (int)j, w = gp.rec_word[j];
                if (packed[(size_t)w]) continue;
                packed[(size_t)w] = true;
                snprintf(b, sizeof b, "        rec[%d] = %s;\n", w, to_word(k).c_str()); s += b;
                if (bp.proj[k].nullable) { snprintf(b, sizeof b, "        m |= o.n%d ? (1ull << %d) : 0ull;\n", k, 40 + w - 1 - nk); s += b; }
            }
            s += "        rec[0] = m;\n    }\n";
            s += "    static __device__ __forceinline__ void record_unpack(const qe_u64 (&rec)[RW], Out &o) {\n";
            for (size_t j = 0; j < bp.agg_fns.size(); j++) {
                const int k = gp.nkeys + (int)j, w = gp.rec_word[j];
                snprintf(b, sizeof b, "        o.o%d = %s;\n", k, from_word(k, w).c_str()); s += b;
                if (bp.proj[k].nullable) { snprintf(b, sizeof b, "        o.n%d = (rec[0] >> %d) & 1ull;\n", k, 40 + w - 1 - nk); s += b; }
            }
            s += "    }\n";
        }
        // every SUM / AVG input of the row is finite (whether null or not: a null's value bits are arbitrary)
        s += "    static __device__ __forceinline__ bool all_finite(const Out &o) {\n        bool f = true;\n";
        {
            std::vector<bool> seen((size_t)gp.nwords, false);
            for (size_t j = 0; j < bp.agg_fns.size(); j++) {
                const int fn = bp.agg_fns[j], vw = gp.val_word[j];
                if ((fn != 2 && fn != 4) || vw < 0 || seen[(size_t)vw]) continue;
                seen[(size_t)vw] = true;
                s += "        f = f && (((qe_u32)(qe_f64_bits(" + agg_val((int)j) + ") >> 32) & 0x7ff00000u) != 0x7ff00000u);\n";
            }
        }
        s += "        return f;\n    }\n";
    }
    s += "};\n\n";
    const char *body = mode == KM_FILTER ? (shape.ring ? "qe_filter_ring_body" : shape.staged ? "qe_filter_staged_body" : shape.pipelined ? "qe_filter_pipelined_body" : "qe_filter_body")
                     : mode == KM_PROJECT ? (shape.pring ? "qe_project_ring_body" : "qe_project_body") : mode == KM_AGG ? "qe_agg_body"
                     : mode == KM_GROUPBY_PRIVATE ? "qe_groupby_private_body"
                     : mode == KM_GROUPBY_REG ? "qe_groupby_reg_body"
                     : mode == KM_GROUPBY_SHARED ? (shape.gpass == 1 ? "qe_groupby_scatter_body" : shape.gpass == 2 ? "qe_groupby_gather_body"
                                                    : shape.scopies < 0 ? "qe_groupby_shared_body" : "qe_groupby_copies_body") : "qe_groupby_atomic_body";
    snprintf(b, sizeof b,
             "extern \"C\" __global__ void __launch_bounds__(%d, %d) %s"
             "(const __grid_constant__ QeParams p)\n{\n    %s<P>(p);\n}\n",
             (mode == KM_FILTER || mode == KM_GROUPBY_REG || (mode == KM_PROJECT && shape.pring)) ? shape.block_threads()
                 : (mode == KM_GROUPBY_PRIVATE || mode == KM_GROUPBY_SHARED) ? shape.gthreads : 256, shape.min_blocks,
             jit_entry_name(mode, shape), body);
    s += b;
    return s;
}

const char *jit_entry_name(KernelMode mode, const TileShape &shape)
{
    switch (mode) {
    case KM_FILTER: return shape.ring ? "qe_filter_ring_kernel" : shape.staged ? "qe_filter_staged_kernel"
                         : shape.pipelined ? "qe_filter_pipelined_kernel" : "qe_filter_kernel";
    case KM_PROJECT: return shape.pring ? "qe_project_ring_kernel" : "qe_project_kernel";
    case KM_AGG: return "qe_agg_kernel";
    case KM_GROUPBY_PRIVATE: return "qe_groupby_private_kernel";
    case KM_GROUPBY_SHARED: return shape.gpass == 1 ? "qe_groupby_scatter_kernel" : shape.gpass == 2 ? "qe_groupby_gather_kernel"
                                 : shape.scopies == 32 ? "qe_groupby_lane_kernel" : shape.scopies == 16 ? "qe_groupby_lane16_kernel"
                                 : shape.scopies == 8 ? "qe_groupby_lane8_kernel" : shape.scopies == 4 ? "qe_groupby_lane4_kernel"
                                 : shape.scopies == 2 ? "qe_groupby_lane2_kernel"
                                 : shape.scopies < 0 ? "qe_groupby_warp_kernel" : "qe_groupby_shared_kernel";
    case KM_GROUPBY_REG: return "qe_groupby_reg_kernel";
    default: return "qe_groupby_atomic_kernel";
    }
}

}  // namespace qe
