// qe_interp.cuh -- generic pipeline program for the kernel bodies of
// qe_device.cuh: a per-row stack interpreter over the serialized pipeline in
// QeParams::iprog.  It is the always-available GPU path (no NVRTC needed) and
// the cross-check for generated code; generated code is what the benchmarks run.
//
// Same semantics as the generated code: every opcode calls the same functions
// of qe_ops.h (reference: M/evaluator/Interpreter.kt:29-109).
//
// iprog layout (32-bit words), written by qe_runtime.cu:
//   [0] has_filter  [1] nproj  [2] nagg  [3] nslots
//   [4  + s]  slot s:  type | nullable << 8                      (16 entries)
//   [20 + k]  out k:   type | has_validity << 8 | (identity_slot + 1) << 16   (8 entries)
//   [28 + k]  aggregate function of projection k                 (8 entries)
//   [36 ...]  programs: filter (if any) then projections, each  len, word[len]
//             word = opcode | arg << 8 ; CONST_F64 is followed by two payload words, CONST_STR by one
// STRING values (dictionary codes, int32) travel on the double stack: the widening is exact and
// equality of the doubles is equality of the codes.
#ifndef QE_INTERP_CUH
#define QE_INTERP_CUH

#include "qe_device.cuh"

#define QE_IP_SLOTS 4
#define QE_IP_OUTS 20
#define QE_IP_AGGS 28
#define QE_IP_PROGRAMS 36
#define QE_ISTACK 32

struct QeInterpValue {
    double f;
    bool b;
    bool valid;
};

// Evaluate one program for table row r.  Payloads: doubles on a small local
// stack, booleans and validity as bit masks indexed by stack depth.
static __device__ __noinline__ QeInterpValue qe_interp_run(const QeParams &p, const qe_u32 *code, int len, qe_i64 r)
{
    double sf[QE_ISTACK];
    qe_u32 sb = 0, sv = 0;      // bit i: boolean payload / validity of stack entry i
    int sp = 0;
    const qe_u32 *hdr = p.iprog;
    for (int pc = 0; pc < len; pc++) {
        const qe_u32 w = code[pc];
        const qe_u32 op = w & 0xffu, arg = w >> 8;
        if (op == 32u) {                                        // LOAD_COL
            const qe_u32 info = hdr[QE_IP_SLOTS + arg];
            const qe_u32 type = info & 0xffu;
            bool valid = true;
            if (info >> 8) valid = (p.col_valid[arg][r >> 5] >> ((qe_u32)r & 31u)) & 1u;
            double f = 0.0;
            bool bv = false;
            const void *col = p.col_values[arg];
            if (type == QE_T_F64) f = ((const double *)col)[r];
            else if (type == QE_T_I64) f = qe_i64_to_f64(((const qe_i64 *)col)[r]);
            else if (type == QE_T_DATE32 || type == QE_T_STR) f = qe_i32_to_f64(((const qe_i32 *)col)[r]);
            else bv = (((const qe_u32 *)col)[r >> 5] >> ((qe_u32)r & 31u)) & 1u;
            sf[sp] = f;
            sb = (sb & ~(1u << sp)) | ((bv ? 1u : 0u) << sp);
            sv = (sv & ~(1u << sp)) | ((valid ? 1u : 0u) << sp);
            sp++;
        } else if (op == 33u) {                                 // CONST_F64
            const qe_u64 bits = (qe_u64)code[pc + 1] | ((qe_u64)code[pc + 2] << 32);
            pc += 2;
            sf[sp] = qe_bits_f64((qe_i64)bits);
            sv |= 1u << sp;
            sp++;
        } else if (op == 35u) {                                 // CONST_STR
            pc += 1;
            sf[sp] = qe_i32_to_f64((qe_i32)code[pc]);
            sv |= 1u << sp;
            sp++;
        } else if (op == 34u) {                                 // CONST_BOOL
            sb = (sb & ~(1u << sp)) | ((arg & 1u) << sp);
            sv |= 1u << sp;
            sp++;
        } else if (op == 3u || op == 4u || op == 5u) {          // NOT, UNARY_MINUS, UNARY_PLUS
            const int a = sp - 1;
            if (op == 3u) sb ^= 1u << a;
            if (op == 4u) sf[a] = qe_neg(sf[a]);
        } else if (op == 2u) {                                  // IF(c, t, e)
            const int c = sp - 3, t = sp - 2, e = sp - 1;
            const bool cv = (sv >> c) & 1u, cb = (sb >> c) & 1u;
            const int pick = cb ? t : e;
            const double f = sf[pick];
            const bool bv = (sb >> pick) & 1u;
            const bool valid = cv && ((sv >> pick) & 1u);
            sf[c] = f;
            sb = (sb & ~(1u << c)) | ((bv ? 1u : 0u) << c);
            sv = (sv & ~(1u << c)) | ((valid ? 1u : 0u) << c);
            sp -= 2;
        } else {                                                // binary
            const int a = sp - 2, b = sp - 1;
            const bool va = (sv >> a) & 1u, vb = (sv >> b) & 1u;
            const bool ba = (sb >> a) & 1u, bb = (sb >> b) & 1u;
            const double fa = sf[a], fb = sf[b];
            bool valid = va && vb, bv = false;
            double f = 0.0;
            // bit 0 of arg: CMP_EQ / CMP_NE of two booleans (both payload doubles are then unused; set when the
            // pipeline is bound), or QE_CMP_IEEE on an ordering comparison (set by the caller)
            const bool bool_cmp = arg & 1u, ieee = arg & 1u;
            switch (op) {
            case 0u: qe_and(ba, va, bb, vb, bv, valid); break;
            case 1u: qe_or(ba, va, bb, vb, bv, valid); break;
            case 6u: f = qe_mul(fa, fb); break;
            case 7u: f = qe_div(fa, fb); break;
            case 8u: f = qe_mod(fa, fb); break;
            case 9u: f = qe_add(fa, fb); break;
            case 10u: f = qe_sub(fa, fb); break;
            case 11u: bv = ieee ? fa < fb : qe_cmp_lt(fa, fb); break;      // ClosureCompiler.kt:127-130 / Interpreter.kt:104-107
            case 12u: bv = ieee ? fa <= fb : qe_cmp_le(fa, fb); break;
            case 13u: bv = ieee ? fa >= fb : qe_cmp_ge(fa, fb); break;
            case 14u: bv = ieee ? fa > fb : qe_cmp_gt(fa, fb); break;
            case 15u: bv = bool_cmp ? (ba == bb) : qe_cmp_eq(fa, fb); break;
            case 16u: bv = bool_cmp ? (ba != bb) : qe_cmp_ne(fa, fb); break;
            }
            sf[a] = f;
            sb = (sb & ~(1u << a)) | ((bv ? 1u : 0u) << a);
            sv = (sv & ~(1u << a)) | ((valid ? 1u : 0u) << a);
            sp -= 1;
        }
    }
    QeInterpValue out;
    out.f = sf[0];
    out.b = sb & 1u;
    out.valid = sv & 1u;
    return out;
}

template <int STEPS_, int VEC_>
struct QeInterpProgram {
    static constexpr int STEPS = STEPS_, VEC = VEC_, NBITCOLS = 2 * QE_MAX_OUTS;

    struct Row { qe_i64 r; };
    struct Out {
        double f[QE_MAX_OUTS];     // DOUBLE payloads, or raw bits of identity INT64 / DATE32 columns
        qe_u32 b;                  // BOOLEAN payloads, bit k
        qe_u32 v;                  // validity, bit k
    };

    static __device__ __forceinline__ void load(const QeParams &p, qe_i64 r, Row (&rows)[VEC])
    {
#pragma unroll
        for (int j = 0; j < VEC; j++) rows[j].r = r + j;      // columns are read on demand by LOAD_COL
    }

    static __device__ bool eval(const QeParams &p, const Row &in, Out &o)
    {
        const qe_u32 *hdr = p.iprog;
        const qe_u32 *code = hdr + QE_IP_PROGRAMS;
        o.b = 0;
        o.v = 0;
        if (in.r >= p.nrows) return false;                    // padded rows: nothing to evaluate
        bool keep = true;
        if (hdr[0]) {
            const int len = (int)code[0];
            const QeInterpValue f = qe_interp_run(p, code + 1, len, in.r);
            keep = f.valid && f.b;                            // M/operator/FilterOperator.kt:20
            code += 1 + len;
        }
        const int nproj = (int)hdr[1];
        for (int k = 0; k < nproj; k++) {
            const int len = (int)code[0];
            const qe_u32 oi = hdr[QE_IP_OUTS + k];
            const qe_u32 type = oi & 0xffu;
            const QeInterpValue v = qe_interp_run(p, code + 1, len, in.r);
            code += 1 + len;
            double f = v.f;
            const int ident = (int)(oi >> 16) - 1;
            if (ident >= 0 && type == QE_T_I64) f = qe_bits_f64(((const qe_i64 *)p.col_values[ident])[in.r]);
            if (ident >= 0 && type == QE_T_DATE32) f = qe_bits_f64((qe_i64)((const qe_i32 *)p.col_values[ident])[in.r]);
            o.f[k] = f;
            o.b |= ((v.b && v.valid) ? 1u : 0u) << k;
            o.v |= (v.valid ? 1u : 0u) << k;
        }
        return keep;
    }

    static __device__ void store_value(const QeParams &p, int k, qe_u64 pos, double f)
    {
        const qe_u32 type = p.iprog[QE_IP_OUTS + k] & 0xffu;
        if (type == QE_T_F64) ((double *)p.out_values[k])[pos] = f;
        else if (type == QE_T_I64) ((qe_i64 *)p.out_values[k])[pos] = qe_f64_bits(f);
        else if (type == QE_T_DATE32) ((qe_i32 *)p.out_values[k])[pos] = (qe_i32)qe_f64_bits(f);
        else if (type == QE_T_STR) ((qe_i32 *)p.out_values[k])[pos] = (qe_i32)f;
    }

    static __device__ void emit(const QeParams &p, qe_u64 pos, const Out &o)
    {
        const int nproj = (int)p.iprog[1];
        for (int k = 0; k < nproj; k++) store_value(p, k, pos, o.f[k]);
    }

    static __device__ void emit_dense(const QeParams &p, qe_i64 r, const Out (&o)[VEC], qe_u32 inb)
    {
        const int nproj = (int)p.iprog[1];
        for (int k = 0; k < nproj; k++)
            for (int j = 0; j < VEC; j++)
                if ((inb >> j) & 1u) store_value(p, k, (qe_u64)(r + j), o[j].f[k]);
    }

    // bit column b < QE_MAX_OUTS: validity of output b; otherwise BOOLEAN payload of output b - QE_MAX_OUTS
    static __device__ __forceinline__ qe_u32 *bitcol(const QeParams &p, int b)
    {
        if (b < QE_MAX_OUTS) return b < (int)p.iprog[1] ? p.out_valid[b] : 0;
        const int k = b - QE_MAX_OUTS;
        if (k >= (int)p.iprog[1]) return 0;
        return (p.iprog[QE_IP_OUTS + k] & 0xffu) == QE_T_BOOL ? (qe_u32 *)p.out_values[k] : 0;
    }

    static __device__ __forceinline__ bool bit(const Out &o, int b)
    {
        return b < QE_MAX_OUTS ? (o.v >> b) & 1u : (o.b >> (b - QE_MAX_OUTS)) & 1u;
    }

    struct Accs { QeAcc a[QE_MAX_AGGS]; };

    static __device__ void acc_init(Accs &a)
    {
        for (int k = 0; k < QE_MAX_AGGS; k++) qe_acc_init(a.a[k]);
    }

    static __device__ void acc_add(const QeParams &p, Accs &a, const Out &o, bool keep)
    {
        const int nagg = (int)p.iprog[2];
        for (int k = 0; k < nagg; k++) {
            const qe_u32 type = p.iprog[QE_IP_OUTS + k] & 0xffu;
            double v = 0.0;
            if (type == QE_T_F64) v = o.f[k];
            else if (type == QE_T_I64 || type == QE_T_DATE32) v = qe_i64_to_f64(qe_f64_bits(o.f[k]));
            if (keep && ((o.v >> k) & 1u)) {
                QeAcc one;
                one.count = 1; one.sum = v; one.mn = v; one.mx = v;
                qe_acc_merge_all(a.a[k], one);
            }
        }
    }

    static __device__ void acc_store(const QeParams &p, Accs &a, QeAcc *slot, qe_i64 part)
    {
        const int nagg = (int)p.iprog[2];
        for (int k = 0; k < nagg; k++) qe_acc_block_reduce_store_all(a.a[k], slot, p.agg_partials, part, nagg, k);
    }
};

#endif  // QE_INTERP_CUH
