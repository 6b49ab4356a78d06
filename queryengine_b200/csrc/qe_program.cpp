// qe_program.cpp -- see qe_program.h
#include "qe_program.h"

#include <cmath>
#include <cstdio>
#include <cstring>

namespace qe {

static const char *const kOpNames[17] = {
    "AND", "OR", "IF", "NOT", "UNARY_MINUS", "UNARY_PLUS", "MUL", "DIV", "MOD", "ADD", "SUB",
    "CMP_LT", "CMP_LE", "CMP_GE", "CMP_GT", "CMP_EQ", "CMP_NE"};

const char *op_name(int op)
{
    if (op >= 0 && op < 17) return kOpNames[op];
    if (op == OP_LOAD_COL) return "LOAD_COL";
    if (op == OP_CONST_F64) return "CONST_F64";
    if (op == OP_CONST_BOOL) return "CONST_BOOL";
    if (op == OP_CONST_STR) return "CONST_STR";
    return "?";
}

// arity of each Function, M/ast/Functions.kt:7-22
int op_arity(int op)
{
    switch (op) {
    case OP_IF: return 3;
    case OP_NOT: case OP_UNARY_MINUS: case OP_UNARY_PLUS: return 1;
    case OP_LOAD_COL: case OP_CONST_F64: case OP_CONST_BOOL: case OP_CONST_STR: return 0;
    default: return (op >= 0 && op < 17) ? 2 : -1;
    }
}

bool parse_program(const uint32_t *words, size_t n, Program &out, std::string &err)
{
    out = Program();
    if (!words || n == 0) { err = "empty program"; return false; }
    if (n > 1024) { err = "program longer than QE_MAX_PROGRAM_WORDS"; return false; }
    out.words.assign(words, words + n);
    int depth = 0;
    for (size_t pc = 0; pc < n; pc++) {
        Instr in;
        in.op = (uint8_t)(words[pc] & 0xffu);
        in.arg = words[pc] >> 8;
        const int ar = op_arity(in.op);
        if (ar < 0) {
            char b[64]; snprintf(b, sizeof b, "unknown opcode %u at word %zu", in.op, pc);
            err = b; return false;
        }
        if (in.op == OP_CONST_F64) {
            if (pc + 2 >= n) { err = "CONST_F64 truncated"; return false; }
            uint64_t bits = (uint64_t)words[pc + 1] | ((uint64_t)words[pc + 2] << 32);
            memcpy(&in.k, &bits, sizeof bits);
            pc += 2;
        }
        if (in.op == OP_CONST_STR) {
            if (pc + 1 >= n) { err = "CONST_STR truncated"; return false; }
            in.code = (int32_t)words[pc + 1];
            pc += 1;
        }
        if (in.op == OP_CONST_BOOL && in.arg > 1) { err = "CONST_BOOL argument must be 0 or 1"; return false; }
        // ordering comparisons: argument bit 0 = QE_CMP_IEEE (the CLOSURE_COMPILER mode's unboxed `<`, ClosureCompiler.kt:127-130)
        if (in.op >= OP_CMP_LT && in.op <= OP_CMP_GT && in.arg > 1) { err = "comparison argument must be 0 or QE_CMP_IEEE"; return false; }
        if (ar > 0 && !(in.op >= OP_CMP_LT && in.op <= OP_CMP_GT) && in.arg != 0) { err = std::string(op_name(in.op)) + " takes no argument"; return false; }
        if (depth < ar) {
            err = std::string("stack underflow at ") + op_name(in.op); return false;
        }
        depth += 1 - ar;
        if (depth > out.max_stack) out.max_stack = depth;
        if (depth > 32) { err = "expression deeper than QE_MAX_STACK"; return false; }
        out.code.push_back(in);
    }
    if (depth != 1) { err = "program must leave exactly one value"; return false; }
    return true;
}

namespace {

// type and nullability of a stack entry, plus what the code generator's special cases look at: is it a literal
// (index of its CONST instruction) or a bare INT64 / DATE32 column
struct TV { VType t; bool nullable; int const_at = -1; bool raw_int = false; };

const char *vt_name(VType t) { return t == VT_F64 ? "DOUBLE" : t == VT_BOOL ? "BOOLEAN" : t == VT_STR ? "STRING" : "?"; }

std::string invalid_types(int op, const TV *a, int n)
{
    // TypeCheck.invalidTypes, M/evaluator/TypeCheck.kt:123-125
    std::string s = std::string("Invalid operand types for [") + op_name(op) + "] [";
    for (int i = 0; i < n; i++) { if (i) s += ", "; s += vt_name(a[i].t); }
    return s + "]";
}

VType value_type_of_column(int qtype)
{
    switch (qtype) {
    case 1: case 3: case 4: return VT_F64;    // DOUBLE, INT64, DATE32 (widened on load)
    case 2: return VT_BOOL;
    case 0: return VT_STR;                    // dictionary codes
    default: return VT_NONE;
    }
}

}  // namespace

static int bind_expr(const std::vector<ColumnInfo> &schema, const Program &p, BoundPipeline &bp,
                     BoundExpr &out, std::string &err)
{
    out = BoundExpr();
    std::vector<TV> st;
    for (const Instr &src : p.code) {
        Instr in = src;
        switch (in.op) {
        case OP_LOAD_COL: {
            if (in.arg >= schema.size()) { err = "column index out of range"; return 2; }
            const ColumnInfo &ci = schema[in.arg];
            const VType t = value_type_of_column(ci.type);
            if (t == VT_NONE) { err = "unknown column type"; return 1; }
            int slot = -1;
            for (size_t s = 0; s < bp.slot_table_col.size(); s++)
                if (bp.slot_table_col[s] == (int)in.arg) slot = (int)s;
            if (slot < 0) {
                if (bp.slot_table_col.size() >= 16) { err = "pipeline references more than QE_MAX_PIPELINE_COLUMNS columns"; return 4; }
                slot = (int)bp.slot_table_col.size();
                bp.slot_table_col.push_back((int)in.arg);
                bp.slot_info.push_back(ci);
            }
            in.arg = (uint32_t)slot;
            st.push_back({t, ci.nullable, -1, ci.type == 3 || ci.type == 4});
            break;
        }
        case OP_CONST_F64: st.push_back({VT_F64, false, (int)out.code.size(), false}); break;
        case OP_CONST_BOOL: st.push_back({VT_BOOL, false}); break;
        case OP_CONST_STR: st.push_back({VT_STR, false}); break;
        default: {
            const int ar = op_arity(in.op);
            TV a[3];
            for (int i = ar - 1; i >= 0; i--) { a[i] = st.back(); st.pop_back(); }
            TV r{VT_NONE, false};
            switch (in.op) {
            case OP_UNARY_MINUS: case OP_UNARY_PLUS:                      // TypeCheck.kt:50-57
                if (a[0].t != VT_F64) { err = invalid_types(in.op, a, ar); return 2; }
                r = {VT_F64, a[0].nullable};
                break;
            case OP_ADD: case OP_SUB: case OP_MUL: case OP_DIV: case OP_MOD:
                if (a[0].t != VT_F64 || a[1].t != VT_F64) { err = invalid_types(in.op, a, ar); return 2; }
                r = {VT_F64, a[0].nullable || a[1].nullable};
                if (in.op == OP_MOD && a[1].const_at >= 0 && a[0].const_at < 0) out.code[(size_t)a[1].const_at].lit_need |= LIT_MOD;
                break;
            case OP_NOT:                                                  // TypeCheck.kt:58-64
                if (a[0].t != VT_BOOL) { err = invalid_types(in.op, a, ar); return 2; }
                r = {VT_BOOL, a[0].nullable};
                break;
            case OP_CMP_EQ: case OP_CMP_NE:                               // TypeCheck.kt:65-71
                if (a[0].t != a[1].t) { err = invalid_types(in.op, a, ar); return 2; }
                for (int i = 0; i < 2 && a[0].t == VT_F64; i++)
                    if (a[i].const_at >= 0) out.code[(size_t)a[i].const_at].lit_need |= LIT_PLAIN | (a[1 - i].raw_int ? LIT_INT : 0u);
                in.arg = a[0].t == VT_BOOL ? 1u : 0u;     // Boolean equality (read by the interpreter kernels); STRING
                                                          // codes compare like numbers
                r = {VT_BOOL, a[0].nullable || a[1].nullable};
                break;
            case OP_CMP_LT: case OP_CMP_LE: case OP_CMP_GE: case OP_CMP_GT:   // TypeCheck.kt:72-78
                if (a[0].t != VT_F64 || a[1].t != VT_F64) { err = invalid_types(in.op, a, ar); return 2; }
                for (int i = 0; i < 2; i++)
                    if (a[i].const_at >= 0) out.code[(size_t)a[i].const_at].lit_need |= LIT_PLAIN | (a[1 - i].raw_int ? LIT_INT : 0u);
                r = {VT_BOOL, a[0].nullable || a[1].nullable};
                break;
            case OP_AND: case OP_OR:                                      // TypeCheck.kt:79-85, intent
                if (a[0].t != VT_BOOL || a[1].t != VT_BOOL) { err = invalid_types(in.op, a, ar); return 2; }
                r = {VT_BOOL, a[0].nullable || a[1].nullable};
                break;
            case OP_IF:                                                   // TypeCheck.kt:86-94
                if (a[0].t != VT_BOOL || a[1].t != a[2].t) { err = invalid_types(in.op, a, ar); return 2; }
                r = {a[1].t, a[0].nullable || a[1].nullable || a[2].nullable};
                break;
            }
            st.push_back(r);
        }
        }
        out.code.push_back(in);
    }
    out.type = st.back().t;
    out.nullable = st.back().nullable;
    if (out.code.size() == 1 && out.code[0].op == OP_LOAD_COL) out.identity_slot = (int)out.code[0].arg;
    return 0;
}

bool Literal::plain() const { return !is_str && k == k && k != 0.0; }
bool Literal::int_ok() const { return !is_str && k == k && std::fabs(k) < 4.0e18 && !(k == 0.0 && std::signbit(k)); }
bool Literal::integral() const { return int_ok() && std::floor(k) == std::ceil(k); }
bool Literal::mod_ok() const
{
    if (is_str || !(k == k)) return false;
    const double kk = std::fabs(k);
    return kk >= 0x1p-1000 && kk <= 0x1p1000 && std::isfinite(1.0 / kk);
}

// Number the distinct literals of the pipeline in order of first appearance.  Equal values share a slot, so
// repeated subtrees still print identically and the code generator's common-subexpression table finds them.
static void assign_literals(BoundPipeline &bp)
{
    auto visit = [&](BoundExpr &e) {
        for (Instr &in : e.code) {
            if (in.op != OP_CONST_F64 && in.op != OP_CONST_STR) continue;
            const bool str = in.op == OP_CONST_STR;
            uint64_t bits;
            memcpy(&bits, &in.k, sizeof bits);
            int slot = -1;
            for (size_t i = 0; i < bp.lits.size() && slot < 0; i++) {
                const Literal &l = bp.lits[i];
                uint64_t lb;
                memcpy(&lb, &l.k, sizeof lb);
                if (l.is_str == str && (str ? l.code == in.code : lb == bits)) slot = (int)i;
            }
            if (slot < 0 && bp.lits.size() < QE_MAX_LITS) {
                Literal l;
                l.is_str = str;
                l.k = str ? 0.0 : in.k;
                l.code = str ? in.code : 0;
                slot = (int)bp.lits.size();
                bp.lits.push_back(l);
            }
            in.lit = slot;
            if (slot >= 0) bp.lits[(size_t)slot].need |= in.lit_need;
        }
    };
    if (bp.has_filter) visit(bp.filter);
    for (BoundExpr &e : bp.proj) visit(e);
}

int bind_pipeline(const std::vector<ColumnInfo> &schema,
                  const Program *filter, const std::vector<const Program *> &proj,
                  const std::vector<int> &agg_fns, BoundPipeline &out, std::string &err, int group_keys)
{
    out = BoundPipeline();
    out.group_keys = group_keys;
    const size_t nkeys = group_keys > 0 ? (size_t)group_keys : 0;
    if (group_keys >= 0) {
        if (proj.size() > 16) { err = "more than 16 group keys + aggregate inputs"; return 4; }
        if (nkeys > 8 || nkeys > proj.size()) { err = "bad number of group keys"; return 1; }
        if (agg_fns.size() != proj.size() - nkeys) { err = "one aggregate function per projection after the keys"; return 1; }
    } else {
        if (proj.size() > 8) { err = "more than QE_MAX_PROJECTIONS projections"; return 4; }
        if (!agg_fns.empty() && agg_fns.size() != proj.size()) {
            err = "each aggregate consumes one projection: nagg must equal nproj"; return 1;
        }
    }
    if (!filter && proj.empty()) { err = "pipeline has neither filter nor projection"; return 1; }
    if (filter) {
        out.has_filter = true;
        int rc = bind_expr(schema, *filter, out, out.filter, err);
        if (rc) return rc;
        if (out.filter.type != VT_BOOL) { err = "filter expression must be BOOLEAN"; return 2; }
    }
    for (const Program *p : proj) {
        if (!p) { err = "null projection program"; return 1; }
        BoundExpr be;
        int rc = bind_expr(schema, *p, out, be, err);
        if (rc) return rc;
        BoundOutput bo;
        if (be.identity_slot >= 0) bo.type = out.slot_info[be.identity_slot].type;   // bare column keeps its physical type
        else bo.type = be.type == VT_F64 ? 1 : be.type == VT_BOOL ? 2 : 0;
        bo.has_validity = be.nullable;
        out.proj.push_back(be);
        out.outs.push_back(bo);
    }
    for (size_t k = 0; k < agg_fns.size(); k++) {
        const int fn = agg_fns[k];
        if (fn == 5 || fn == 6) { err = "ANY / ALL aggregates are not implemented (as in the reference)"; return 4; }
        if (fn < 0 || fn > 4) { err = "unknown aggregation function"; return 1; }
        // M/evaluator/TypeCheck.kt:110-120: MIN/MAX/SUM/AVG need DOUBLE, COUNT takes anything
        if (fn != 3 && out.proj[nkeys + k].type != VT_F64) {
            err = "Invalid operand types for aggregation: DOUBLE expected"; return 2;
        }
        out.agg_fns.push_back(fn);
    }
    assign_literals(out);
    return 0;
}

std::string BoundPipeline::signature() const
{
    std::string s;
    char b[64];
    auto put_expr = [&](const BoundExpr &e) {
        for (const Instr &in : e.code) {
            uint64_t kb;
            memcpy(&kb, &in.k, sizeof kb);
            if ((in.op == OP_CONST_F64 || in.op == OP_CONST_STR) && in.lit >= 0)      // a parameter: slot + class, not the value
                snprintf(b, sizeof b, "%u.L%d.%x,", in.op, in.lit, lits[(size_t)in.lit].flags());
            else
                snprintf(b, sizeof b, "%u.%u.%llx,", in.op, in.arg,
                         (unsigned long long)(in.op == OP_CONST_F64 ? kb : in.op == OP_CONST_STR ? (uint64_t)(uint32_t)in.code : 0));
            s += b;
        }
        s += ';';
    };
    s += "C";
    for (const ColumnInfo &c : slot_info) { snprintf(b, sizeof b, "%d%c", c.type, c.nullable ? 'n' : '.'); s += b; }
    s += has_filter ? "|F" : "|-";
    if (has_filter) put_expr(filter);
    s += "|P";
    for (const BoundExpr &e : proj) put_expr(e);
    s += "|A";
    for (int f : agg_fns) { snprintf(b, sizeof b, "%d,", f); s += b; }
    if (group_keys >= 0) { snprintf(b, sizeof b, "|G%d", group_keys); s += b; }
    return s;
}

}  // namespace qe
