// qe_program.h -- host-side representation of flattened expression programs:
// decoding, structural validation, type checking against a table schema and
// binding of a whole pipeline (filter + projections + aggregates) to column
// slots.  Pure host C++, no CUDA.
//
// The type rules restate M/evaluator/TypeCheck.kt:38-96 (M/ = the reference's
// src/main/java/net/jhorstmann/queryengine/), with the evident intent for AND/OR
// (BOOLEAN, BOOLEAN): the reference's rule at TypeCheck.kt:80 demands a DOUBLE
// left operand, which rejects every real predicate.
#ifndef QE_PROGRAM_H
#define QE_PROGRAM_H

#include <cstdint>
#include <string>
#include <vector>

namespace qe {

// opcodes 0..16 = reference Function ordinals (M/ast/Functions.kt:7-22)
enum Op : uint8_t {
    OP_AND = 0, OP_OR, OP_IF, OP_NOT, OP_UNARY_MINUS, OP_UNARY_PLUS,
    OP_MUL, OP_DIV, OP_MOD, OP_ADD, OP_SUB,
    OP_CMP_LT, OP_CMP_LE, OP_CMP_GE, OP_CMP_GT, OP_CMP_EQ, OP_CMP_NE,
    OP_LOAD_COL = 32, OP_CONST_F64 = 33, OP_CONST_BOOL = 34, OP_CONST_STR = 35
};

const char *op_name(int op);
int op_arity(int op);          // -1 for unknown opcodes

// VT_F64 / VT_BOOL == reference DataType ordinals; VT_STR is a STRING held as its dictionary code (int32):
// equal codes <=> equal strings, as long as one dictionary encodes every STRING column and literal of a context
enum VType : uint8_t { VT_NONE = 0, VT_F64 = 1, VT_BOOL = 2, VT_STR = 3 };

struct Instr {
    uint8_t op = 0;
    uint32_t arg = 0;          // LOAD_COL: column (table index before binding, slot after); CONST_BOOL: 0/1
    int32_t code = 0;          // CONST_STR: dictionary code of the StringLiteralExpression
    double k = 0.0;            // CONST_F64
    int lit = -1;              // CONST_F64 / CONST_STR after binding: slot of BoundPipeline::lits (-1: embedded in the code)
    unsigned lit_need = 0;     // CONST_F64: which class flags of the literal the code generator looks at for THIS use
};

// A literal of a bound pipeline.  Generated kernels read literal VALUES from the kernel parameters
// (QeParams::lit_*), so `col < 10` and `col < 11` are one kernel; only the literal's CLASS -- the handful of
// properties the code generator branches on -- is part of the pipeline signature.
#define QE_MAX_LITS 24
struct Literal {
    bool is_str = false;
    double k = 0.0;            // NumericLiteralExpression
    int32_t code = 0;          // StringLiteralExpression: dictionary code
    bool plain() const;        // neither NaN nor a zero: Double.compare agrees with the IEEE comparison (qe_cmp_*_k)
    bool int_ok() const;       // usable against a bare INT64 / DATE32 column in the integer domain (finite, in range, not -0.0)
    bool integral() const;     // ... and floor(k) == ceil(k)
    bool mod_ok() const;       // |k| and 1 / |k| are normal: MOD by reciprocal multiply (qe_mod_k)
    unsigned need = 0;         // union of the uses' lit_need: only these flags are part of the signature
    unsigned flags() const { return is_str ? 16u : ((plain() ? 1u : 0u) | (int_ok() ? 2u : 0u) | (integral() ? 4u : 0u) | (mod_ok() ? 8u : 0u)) & need; }
};
enum { LIT_PLAIN = 1u, LIT_INT = 2u | 4u, LIT_MOD = 8u };

struct Program {
    std::vector<uint32_t> words;
    std::vector<Instr> code;
    int max_stack = 0;
};

// Decode + structural checks (known opcodes, stack discipline, exactly one result).
bool parse_program(const uint32_t *words, size_t n, Program &out, std::string &err);

struct ColumnInfo {
    int type = 0;              // qe_type
    bool nullable = false;
};

struct BoundExpr {
    std::vector<Instr> code;   // LOAD_COL args are pipeline slots
    VType type = VT_NONE;
    bool nullable = false;     // can the result be null?
    int identity_slot = -1;    // >= 0: the expression is a bare ColumnExpression of that slot
};

struct BoundOutput {
    int type = 0;              // physical qe_type of the output column
    bool has_validity = false;
};

struct BoundPipeline {
    std::vector<int> slot_table_col;       // slot -> table column
    std::vector<ColumnInfo> slot_info;
    bool has_filter = false;
    BoundExpr filter;
    std::vector<BoundExpr> proj;
    std::vector<BoundOutput> outs;         // one per projection
    std::vector<int> agg_fns;              // empty, or one per projection (group-by: one per projection after the keys)
    int group_keys = -1;                   // >= 0: GroupByAggregationOperator, the first group_keys projections are the key
    std::vector<Literal> lits;             // distinct literals in order of first appearance (at most QE_MAX_LITS; the rest stay embedded)
    // everything that changes the generated code, as a cache key
    std::string signature() const;
};

// status: 0 ok, 2 program/type error, 4 unsupported, 1 invalid argument (values of qe_status)
// group_keys >= 0 binds a group-by: proj = keys then aggregate inputs (up to 16 in all), agg_fns one per input
int bind_pipeline(const std::vector<ColumnInfo> &schema,
                  const Program *filter, const std::vector<const Program *> &proj,
                  const std::vector<int> &agg_fns, BoundPipeline &out, std::string &err, int group_keys = -1);

}  // namespace qe
#endif
