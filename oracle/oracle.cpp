// oracle/oracle.cpp — TEST INFRASTRUCTURE: CPU restatement of the reference's scan -> filter -> project ->
// (hash-)aggregate path. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
// may load this; the product (presto_b200/) never does.
//
// PARITY PINNING: the reference is Java and no JVM exists in this image, so the reference itself cannot be run
// (DESIGN.md). This restatement is pinned against the reference's own known-answer tests instead
// (tests/test_oracle_*.py): TestUnscaledDecimal128Arithmetic vectors, TestDecimalSumAggregation state vectors,
// the closed-form sequence KATs of AbstractTestAggregationFunction, TestHashAggregationOperator / TestAggregationOperator
// expected rows, TestGroupByHash group-id assignment.
//
// What it follows (paths under /root/reference/core/trino-main/src/main/java/io/trino/):
//   filter -> SelectedPositions   operator/project/PageFilter.java:27-50, sql/gen/PageFunctionCompiler.java:459-544
//   projections                   sql/gen/PageFunctionCompiler.java:274-365 (evaluated at selected positions only)
//   expression semantics          sql/gen/{And,Or,Between,In,If,Coalesce,IsNull}CodeGenerator.java, type/*Operators.java
//   grouping                      operator/MultiChannelGroupByHash.java:281-342, BigintGroupByHash.java:213-323
//   accumulators                  operator/aggregation/* (see Accumulator classes below)
//   operator protocol             operator/HashAggregationOperator.java:367-518, AggregationOperator.java:120-186
// Evaluation is column-at-a-time over batches of <= 8192 positions (PageProcessor.MAX_BATCH_SIZE,
// operator/project/PageProcessor.java:56) so that the timed CPU baseline is not an interpreter benchmark.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <deque>
#include <vector>

#include "../include/prestogpu.h"
#include "oracle_plan.h"  // the oracle's own plan reader: nothing of presto_b200/csrc is compiled into this library
#include "decimal128.h"

namespace oracle {

struct OracleError : std::runtime_error {
    int status;
    OracleError(int st, const std::string& m) : std::runtime_error(m), status(st) {}
};

// A column of evaluated values for a list of positions.
struct Vec {
    Type type;
    std::vector<int64_t> i;     // boolean/integer/date/bigint/short decimal
    std::vector<double> d;      // double
    std::vector<Dec128> q;      // long decimal (block representation)
    std::vector<std::string> s; // var-width
    std::vector<uint8_t> null;  // empty = no nulls
    size_t size() const { return type.kind == TypeKind::DOUBLE ? d.size() : type.isLongDecimal() ? q.size() : type.isVarWidth() ? s.size() : i.size(); }
    bool isNull(size_t k) const { return !null.empty() && null[k]; }
    void resize(size_t n) {
        if (type.kind == TypeKind::DOUBLE) d.assign(n, 0.0);
        else if (type.isLongDecimal()) q.assign(n, Dec128());
        else if (type.isVarWidth()) s.assign(n, std::string());
        else i.assign(n, 0);
    }
    void setNull(size_t k) {
        if (null.empty()) null.assign(size(), 0);
        null[k] = 1;
    }
};

// ------------------------------------------------------------------------------------------- block access
struct BlockView {
    const pgpu_column* col;
    const pgpu_column* flat;  // the block holding values (dictionary / value block for DICT / RLE)
    Type type;
    int64_t index(int64_t pos) const {
        if (col->encoding == PGPU_ENC_DICTIONARY) return col->ids[pos];
        if (col->encoding == PGPU_ENC_RLE) return 0;
        return pos;
    }
    bool isNull(int64_t pos) const { return flat->nulls && flat->nulls[index(pos)]; }
};

static BlockView viewOf(const pgpu_column& c, const Type& t) {
    BlockView v;
    v.col = &c;
    v.flat = c.encoding == PGPU_ENC_PLAIN ? &c : c.dictionary;
    v.type = t;
    return v;
}

// Block.getLong / getDouble / getSlice at the given positions (spi/block/LongArrayBlock.java:123-129 etc.)
static Vec gather(const BlockView& b, const int32_t* pos, size_t n, int64_t base) {
    Vec v;
    v.type = b.type;
    v.resize(n);
    const pgpu_column* f = b.flat;
    bool anyNull = f->nulls != nullptr;
    if (anyNull) v.null.assign(n, 0);
    for (size_t k = 0; k < n; k++) {
        int64_t ix = b.index(base + pos[k]);
        if (anyNull && f->nulls[ix]) {
            v.null[k] = 1;
            continue;
        }
        switch (b.type.kind) {
            case TypeKind::BOOLEAN: v.i[k] = ((const int8_t*)f->values)[ix] != 0; break;
            case TypeKind::INTEGER:
            case TypeKind::DATE: v.i[k] = ((const int32_t*)f->values)[ix]; break;
            case TypeKind::BIGINT: v.i[k] = ((const int64_t*)f->values)[ix]; break;
            case TypeKind::DOUBLE: v.d[k] = ((const double*)f->values)[ix]; break;
            case TypeKind::DECIMAL:
                if (b.type.isShortDecimal()) v.i[k] = ((const int64_t*)f->values)[ix];
                else {
                    v.q[k].lo = ((const uint64_t*)f->values)[2 * ix];
                    v.q[k].hi = ((const uint64_t*)f->values)[2 * ix + 1];
                }
                break;
            case TypeKind::VARCHAR:
            case TypeKind::VARBINARY: {
                int32_t o0 = f->offsets[ix], o1 = f->offsets[ix + 1];
                v.s[k].assign((const char*)f->values + o0, (size_t)(o1 - o0));
                break;
            }
        }
    }
    bool any = false;
    for (auto x : v.null) any |= x != 0;
    if (!any) v.null.clear();
    return v;
}


// ------------------------------------------------------------------------------------------- expressions
struct EvalCtx {
    const std::vector<BlockView>* cols;
    int64_t base;  // first row of the batch inside the page
};

static inline int64_t addExact(int64_t a, int64_t b, const char* what) {  // Math.addExact, type/BigintOperators.java:47-56
    int64_t r;
    if (__builtin_add_overflow(a, b, &r)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, std::string("bigint ") + what + " overflow");
    return r;
}

static Dec128 decOf(const Vec& v, size_t k) { return v.type.isLongDecimal() ? v.q[k] : unscaledDecimal(v.i[k]); }

static Vec evalExpr(const Expr& e, const EvalCtx& cx, const std::vector<int32_t>& pos);

static Vec constantVec(const Expr& e, size_t n) {
    Vec v;
    v.type = e.type;
    v.resize(n);
    if (e.isNull) {
        v.null.assign(n, 1);
        return v;
    }
    if (e.type.kind == TypeKind::DOUBLE) {
        double d;
        memcpy(&d, &e.i64, 8);
        std::fill(v.d.begin(), v.d.end(), d);
    } else if (e.type.isLongDecimal()) {
        std::fill(v.q.begin(), v.q.end(), fromI128(e.dec));
    } else {
        std::fill(v.i.begin(), v.i.end(), e.i64);
    }
    return v;
}

static void mergeNulls(Vec& out, const Vec& a) {
    if (a.null.empty()) return;
    if (out.null.empty()) out.null.assign(out.size(), 0);
    for (size_t k = 0; k < a.null.size(); k++) out.null[k] |= a.null[k];
}

template <class F>
static Vec compareVec(const Vec& a, const Vec& b, F cmp) {  // cmp(sign) on a three-way result; doubles use IEEE operators
    Vec r;
    r.type = Type::boolean();
    size_t n = a.size();
    r.resize(n);
    mergeNulls(r, a);
    mergeNulls(r, b);
    for (size_t k = 0; k < n; k++) {
        if (r.isNull(k)) continue;
        r.i[k] = cmp(a, b, k);
    }
    return r;
}

static Vec compareOp(Op op, const Vec& a, const Vec& b) {
    const Type& t = a.type;
    if (t.kind == TypeKind::DOUBLE) {  // spi/type/DoubleType.java:157-210 — plain IEEE comparisons; <> is not(==)
        return compareVec(a, b, [op](const Vec& x, const Vec& y, size_t k) -> int64_t {
            double l = x.d[k], r = y.d[k];
            switch (op) {
                case Op::EQ: return l == r;
                case Op::NE: return !(l == r);
                case Op::LT: return l < r;
                case Op::LE: return l <= r;
                case Op::GT: return r < l;
                default: return r <= l;
            }
        });
    }
    if (t.isLongDecimal()) {
        return compareVec(a, b, [op](const Vec& x, const Vec& y, size_t k) -> int64_t {
            int c = compare(x.q[k], y.q[k]);
            switch (op) {
                case Op::EQ: return c == 0;
                case Op::NE: return c != 0;
                case Op::LT: return c < 0;
                case Op::LE: return c <= 0;
                case Op::GT: return c > 0;
                default: return c >= 0;
            }
        });
    }
    return compareVec(a, b, [op](const Vec& x, const Vec& y, size_t k) -> int64_t {
        int64_t l = x.i[k], r = y.i[k];
        switch (op) {
            case Op::EQ: return l == r;
            case Op::NE: return l != r;
            case Op::LT: return l < r;
            case Op::LE: return l <= r;
            case Op::GT: return l > r;
            default: return l >= r;
        }
    });
}

// SQL three-valued AND/OR of two evaluated operands (AndCodeGenerator / OrCodeGenerator)
static Vec logic2(bool isAnd, const Vec& a, const Vec& b) {
    Vec r;
    r.type = Type::boolean();
    size_t n = a.size();
    r.resize(n);
    for (size_t k = 0; k < n; k++) {
        bool an = a.isNull(k), bn = b.isNull(k);
        bool av = a.i[k] != 0, bv = b.i[k] != 0;
        bool decisive = isAnd ? ((!an && !av) || (!bn && !bv)) : ((!an && av) || (!bn && bv));
        if (decisive) r.i[k] = isAnd ? 0 : 1;
        else if (an || bn) r.setNull(k);
        else r.i[k] = isAnd ? 1 : 0;
    }
    return r;
}

static Vec evalLogic(const Expr& e, const EvalCtx& cx, const std::vector<int32_t>& pos) {
    const bool isAnd = e.kind == ExprKind::AND;
    size_t n = pos.size();
    // state per position: 0 undecided-all-true-so-far(AND)/all-false(OR), 1 saw null, 2 decisive
    std::vector<uint8_t> st(n, 0);
    std::vector<int32_t> active = pos;
    std::vector<size_t> activeIdx(n);
    for (size_t k = 0; k < n; k++) activeIdx[k] = k;
    for (auto& arg : e.args) {
        if (active.empty()) break;
        Vec v = evalExpr(*arg, cx, active);  // only rows that are still undecided are evaluated (short circuit)
        std::vector<int32_t> nextPos;
        std::vector<size_t> nextIdx;
        for (size_t k = 0; k < active.size(); k++) {
            size_t o = activeIdx[k];
            if (v.isNull(k)) st[o] = 1;
            else if ((v.i[k] != 0) != isAnd) { st[o] = 2; continue; }
            nextPos.push_back(active[k]);
            nextIdx.push_back(o);
        }
        active.swap(nextPos);
        activeIdx.swap(nextIdx);
    }
    Vec r;
    r.type = Type::boolean();
    r.resize(n);
    for (size_t k = 0; k < n; k++) {
        if (st[k] == 2) r.i[k] = isAnd ? 0 : 1;
        else if (st[k] == 1) r.setNull(k);
        else r.i[k] = isAnd ? 1 : 0;
    }
    return r;
}


static Vec scatterNew(const Type& t, size_t n) {
    Vec r;
    r.type = t;
    r.resize(n);
    return r;
}
static void copyAt(Vec& dst, size_t k, const Vec& src, size_t j) {
    if (src.isNull(j)) { dst.setNull(k); return; }
    if (dst.type.kind == TypeKind::DOUBLE) dst.d[k] = src.d[j];
    else if (dst.type.isLongDecimal()) dst.q[k] = src.q[j];
    else if (dst.type.isVarWidth()) dst.s[k] = src.s[j];
    else dst.i[k] = src.i[j];
}

static Vec evalArith(const Expr& e, const Vec& a, const Vec& b) {
    size_t n = a.size();
    Vec r = scatterNew(e.type, n);
    mergeNulls(r, a);
    mergeNulls(r, b);
    const Type& t = e.type;
    for (size_t k = 0; k < n; k++) {
        if (r.isNull(k)) continue;
        if (t.kind == TypeKind::DOUBLE) {  // type/DoubleOperators.java:61-94
            double x = a.d[k], y = b.d[k];
            switch (e.op) {
                case Op::ADD: r.d[k] = x + y; break;
                case Op::SUBTRACT: r.d[k] = x - y; break;
                case Op::MULTIPLY: r.d[k] = x * y; break;
                case Op::DIVIDE: r.d[k] = x / y; break;
                default: r.d[k] = std::fmod(x, y); break;
            }
        } else if (t.kind == TypeKind::BIGINT) {  // type/BigintOperators.java:45-108
            int64_t x = a.i[k], y = b.i[k], z = 0;
            switch (e.op) {
                case Op::ADD:
                    if (__builtin_add_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "bigint addition overflow");
                    break;
                case Op::SUBTRACT:
                    if (__builtin_sub_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "bigint subtraction overflow");
                    break;
                case Op::MULTIPLY:
                    if (__builtin_mul_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "bigint multiplication overflow");
                    break;
                case Op::DIVIDE:
                    if (y == 0) throw OracleError(PGPU_DIVISION_BY_ZERO, "Division by zero");
                    if (x == INT64_MIN && y == -1) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "bigint division overflow");
                    z = x / y;
                    break;
                default:
                    if (y == 0) throw OracleError(PGPU_DIVISION_BY_ZERO, "Division by zero");
                    z = y == -1 ? 0 : x % y;
                    break;
            }
            r.i[k] = z;
        } else if (t.kind == TypeKind::INTEGER) {  // type/IntegerOperators.java:45-103 (Math.*Exact on int)
            int32_t x = (int32_t)a.i[k], y = (int32_t)b.i[k], z = 0;
            switch (e.op) {
                case Op::ADD:
                    if (__builtin_add_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "integer addition overflow");
                    break;
                case Op::SUBTRACT:
                    if (__builtin_sub_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "integer subtraction overflow");
                    break;
                case Op::MULTIPLY:
                    if (__builtin_mul_overflow(x, y, &z)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "integer multiplication overflow");
                    break;
                case Op::DIVIDE:
                    if (y == 0) throw OracleError(PGPU_DIVISION_BY_ZERO, "Division by zero");
                    // IntegerOperators.java:79-89 divides Java longs: MIN / -1 = 2^31 leaves the INTEGER domain and fails when written
                    // to a block (spi/type/AbstractIntType.java:85-95); reported here, at the division
                    if (x == INT32_MIN && y == -1) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "integer division overflow: value 2147483648 exceeds MAX_INT");
                    z = x / y;
                    break;
                default:
                    if (y == 0) throw OracleError(PGPU_DIVISION_BY_ZERO, "Division by zero");
                    z = y == -1 ? 0 : x % y;
                    break;
            }
            r.i[k] = z;
        } else {  // DECIMAL — type/DecimalOperators.java
            const Type &ta = a.type, &tb = b.type;
            try {
                if (e.op == Op::ADD || e.op == Op::SUBTRACT) {
                    if (t.isShortDecimal()) {  // addShortShortShort :103-106 / subtractShortShortShort :208-211
                        int64_t ar = (int64_t)pow10_i128(t.scale - ta.scale), br = (int64_t)pow10_i128(t.scale - tb.scale);
                        r.i[k] = e.op == Op::ADD ? a.i[k] * ar + b.i[k] * br : a.i[k] * ar - b.i[k] * br;
                    } else {  // internalAddLongLongLong :151-172 / internalSubtractLongLongLong :230-250
                        Dec128 x = rescale(decOf(a, k), t.scale - ta.scale), y = rescale(decOf(b, k), t.scale - tb.scale);
                        Dec128 z = e.op == Op::ADD ? add(x, y) : subtract(x, y);
                        throwIfOverflows(z);
                        r.q[k] = z;
                    }
                } else if (e.op == Op::MULTIPLY) {
                    if (t.isShortDecimal()) r.i[k] = a.i[k] * b.i[k];  // multiplyShortShortShort :281-284
                    else {
                        Dec128 z = multiply(decOf(a, k), decOf(b, k));  // multiply*Long :286-330
                        throwIfOverflows(z);
                        r.q[k] = z;
                    }
                } else {
                    throw OracleError(PGPU_NOT_SUPPORTED, "decimal divide/modulus");
                }
            } catch (const ArithmeticOverflow&) {
                throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
            }
        }
    }
    return r;
}

static Vec evalCast(const Expr& e, const Vec& a) {
    const Type &f = a.type, &t = e.type;
    if (f == t) return a;
    size_t n = a.size();
    Vec r = scatterNew(t, n);
    mergeNulls(r, a);
    for (size_t k = 0; k < n; k++) {
        if (r.isNull(k)) continue;
        if (f.kind == TypeKind::INTEGER && t.kind == TypeKind::BIGINT) r.i[k] = a.i[k];
        else if ((f.kind == TypeKind::INTEGER || f.kind == TypeKind::BIGINT) && t.kind == TypeKind::DOUBLE) r.d[k] = (double)a.i[k];
        else if (f.kind == TypeKind::BIGINT && t.kind == TypeKind::INTEGER) {  // BigintOperators.castToInteger :127-137
            if (a.i[k] != (int64_t)(int32_t)a.i[k]) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Out of range for integer");
            r.i[k] = a.i[k];
        } else if (f.kind == TypeKind::DOUBLE && t.kind == TypeKind::BIGINT) {  // DoubleOperators.castToLong :153-165 (HALF_UP)
            double v = a.d[k];
            if (std::isnan(v) || std::isinf(v)) throw OracleError(PGPU_INVALID_CAST_ARGUMENT, "Unable to cast to bigint");
            double z = std::trunc(v);
            if (std::fabs(v - z) >= 0.5) z += v < 0 ? -1.0 : 1.0;
            if (!(z >= -9223372036854775808.0 && z < 9223372036854775808.0)) throw OracleError(PGPU_INVALID_CAST_ARGUMENT, "Unable to cast to bigint");
            r.i[k] = (int64_t)z;
        } else if (f.kind == TypeKind::BOOLEAN && t.kind == TypeKind::DOUBLE) r.d[k] = a.i[k] ? 1.0 : 0.0;
        else if (f.kind == TypeKind::BOOLEAN) r.i[k] = a.i[k] ? 1 : 0;
        else if (t.kind == TypeKind::BOOLEAN) r.i[k] = f.kind == TypeKind::DOUBLE ? a.d[k] != 0.0 : a.i[k] != 0;
        else if ((f.kind == TypeKind::BIGINT || f.kind == TypeKind::INTEGER) && t.kind == TypeKind::DECIMAL) {  // type/DecimalCasts.java
            i128 v = (i128)a.i[k] * pow10_i128(t.scale);
            i128 lim = pow10_i128(t.precision);
            if (v >= lim || v <= -lim) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Cannot cast to decimal");
            if (t.isShortDecimal()) r.i[k] = (int64_t)v;
            else r.q[k] = fromI128(v);
        } else if (f.kind == TypeKind::DECIMAL && t.kind == TypeKind::DECIMAL) {  // type/DecimalToDecimalCasts.java
            i128 v = f.isShortDecimal() ? (i128)a.i[k] : toI128(a.q[k]);
            if (t.scale >= f.scale) v *= pow10_i128(t.scale - f.scale);
            else {
                i128 p = pow10_i128(f.scale - t.scale);
                bool neg = v < 0;
                u128 m = neg ? (u128)0 - (u128)v : (u128)v;
                u128 qd = m / (u128)p, rem = m % (u128)p;
                if (rem >= (u128)p - rem) qd += 1;
                v = neg ? -(i128)qd : (i128)qd;
            }
            i128 lim = pow10_i128(t.precision);
            if (v >= lim || v <= -lim) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Cannot cast decimal");
            if (t.isShortDecimal()) r.i[k] = (int64_t)v;
            else r.q[k] = fromI128(v);
        } else {
            throw OracleError(PGPU_NOT_SUPPORTED, "cast");
        }
    }
    return r;
}


static Vec evalExpr(const Expr& e, const EvalCtx& cx, const std::vector<int32_t>& pos) {
    size_t n = pos.size();
    switch (e.kind) {
        case ExprKind::INPUT: return gather((*cx.cols)[e.channel], pos.data(), n, cx.base);
        case ExprKind::CONST: return constantVec(e, n);
        case ExprKind::AND:
        case ExprKind::OR: return evalLogic(e, cx, pos);
        case ExprKind::BETWEEN: {
            // BetweenCodeGenerator: value >= min AND value <= max, the value evaluated once. min is an argument of the first comparison
            // (skipped where the value is NULL); max of the second, which AND evaluates only where the first is not FALSE.
            auto evalWhere = [&](const Expr& x, const std::vector<char>& live) {
                std::vector<int32_t> sub;
                std::vector<size_t> at;
                for (size_t k = 0; k < n; k++)
                    if (live[k]) { sub.push_back(pos[k]); at.push_back(k); }
                if (at.size() == n) return evalExpr(x, cx, pos);
                Vec full = scatterNew(x.type, n);
                full.null.assign(n, 1);
                if (!sub.empty()) {
                    Vec part = evalExpr(x, cx, sub);
                    for (size_t j = 0; j < at.size(); j++) {
                        if (part.isNull(j)) continue;
                        full.null[at[j]] = 0;
                        copyAt(full, at[j], part, j);
                    }
                }
                return full;
            };
            Vec v = evalExpr(*e.args[0], cx, pos);
            std::vector<char> live(n, 1);
            for (size_t k = 0; k < n; k++) live[k] = !v.isNull(k);
            Vec lo = evalWhere(*e.args[1], live);
            Vec ge = compareOp(Op::GE, v, lo);
            for (size_t k = 0; k < n; k++) live[k] = !v.isNull(k) && !(!ge.isNull(k) && !ge.i[k]);
            Vec hi = evalWhere(*e.args[2], live);
            return logic2(true, ge, compareOp(Op::LE, v, hi));
        }
        case ExprKind::IN: {  // InCodeGenerator: NULL value -> NULL; no match and a NULL member -> NULL
            Vec v = evalExpr(*e.args[0], cx, pos);
            Vec r = scatterNew(Type::boolean(), n);
            bool anyNullMember = false;
            for (size_t j = 1; j < e.args.size(); j++) anyNullMember |= e.args[j]->isNull;
            for (size_t k = 0; k < n; k++) {
                if (v.isNull(k)) { r.setNull(k); continue; }
                bool m = false;
                for (size_t j = 1; j < e.args.size() && !m; j++) {
                    const Expr& c = *e.args[j];
                    if (c.isNull) continue;
                    if (v.type.kind == TypeKind::DOUBLE) {
                        double d;
                        memcpy(&d, &c.i64, 8);
                        m = v.d[k] == d;
                    } else if (v.type.isLongDecimal()) m = compare(v.q[k], fromI128(c.dec)) == 0;
                    else m = v.i[k] == c.i64;
                }
                if (m) r.i[k] = 1;
                else if (anyNullMember) r.setNull(k);
            }
            return r;
        }
        case ExprKind::IS_NULL: {
            Vec v = evalExpr(*e.args[0], cx, pos);
            Vec r = scatterNew(Type::boolean(), n);
            for (size_t k = 0; k < n; k++) r.i[k] = v.isNull(k);
            return r;
        }
        case ExprKind::IF: {  // IfCodeGenerator: NULL condition takes the false branch; branches only run where taken
            Vec c = evalExpr(*e.args[0], cx, pos);
            std::vector<int32_t> pt, pf;
            std::vector<size_t> it, jf;
            for (size_t k = 0; k < n; k++) {
                if (!c.isNull(k) && c.i[k]) { pt.push_back(pos[k]); it.push_back(k); }
                else { pf.push_back(pos[k]); jf.push_back(k); }
            }
            Vec r = scatterNew(e.type, n);
            if (!pt.empty()) {
                Vec a = evalExpr(*e.args[1], cx, pt);
                for (size_t j = 0; j < it.size(); j++) copyAt(r, it[j], a, j);
            }
            if (!pf.empty()) {
                Vec b = evalExpr(*e.args[2], cx, pf);
                for (size_t j = 0; j < jf.size(); j++) copyAt(r, jf[j], b, j);
            }
            return r;
        }
        case ExprKind::COALESCE: {  // CoalesceCodeGenerator: operands evaluated left to right while still NULL
            Vec r = scatterNew(e.type, n);
            r.null.assign(n, 1);
            std::vector<int32_t> active = pos;
            std::vector<size_t> idx(n);
            for (size_t k = 0; k < n; k++) idx[k] = k;
            for (auto& arg : e.args) {
                if (active.empty()) break;
                Vec a = evalExpr(*arg, cx, active);
                std::vector<int32_t> np;
                std::vector<size_t> ni;
                for (size_t j = 0; j < active.size(); j++) {
                    if (a.isNull(j)) { np.push_back(active[j]); ni.push_back(idx[j]); }
                    else { r.null[idx[j]] = 0; copyAt(r, idx[j], a, j); }
                }
                active.swap(np);
                idx.swap(ni);
            }
            bool any = false;
            for (auto x : r.null) any |= x != 0;
            if (!any) r.null.clear();
            return r;
        }
        case ExprKind::NULL_IF: {
            // NullIfCodeGenerator.java:70-113: first NULL -> NULL and the second operand is not evaluated there; equal(first, second)
            // true -> NULL; false or NULL -> first
            Vec a = evalExpr(*e.args[0], cx, pos);
            std::vector<int32_t> sub;
            std::vector<size_t> at;
            for (size_t k = 0; k < n; k++)
                if (!a.isNull(k)) { sub.push_back(pos[k]); at.push_back(k); }
            Vec r = a;
            if (!sub.empty()) {
                Vec b = evalExpr(*e.args[1], cx, sub);
                Vec as = scatterNew(a.type, at.size());
                for (size_t j = 0; j < at.size(); j++) copyAt(as, j, a, at[j]);
                Vec eq = compareOp(Op::EQ, as, b);
                for (size_t j = 0; j < at.size(); j++)
                    if (!eq.isNull(j) && eq.i[j]) r.setNull(at[j]);
            }
            return r;
        }
        case ExprKind::SWITCH: {
            // SwitchCodeGenerator.java:113-159: the value is evaluated once; NULL value -> the else branch; otherwise the first WHEN whose
            // equal(operand, value) is true (a NULL comparison is not a match) gives its result; no match -> else, no else -> NULL.
            // Operands and results are evaluated only at the positions that reach / take them.
            Vec v = evalExpr(*e.args[0], cx, pos);
            Vec r = scatterNew(e.type, n);
            r.null.assign(n, 1);
            std::vector<size_t> open;  // positions (indices into pos) still looking for a match
            std::vector<size_t> toElse;
            for (size_t k = 0; k < n; k++) (v.isNull(k) ? toElse : open).push_back(k);
            auto place = [&](const Expr& x, const std::vector<size_t>& where) {
                if (where.empty()) return;
                std::vector<int32_t> sub;
                for (size_t k : where) sub.push_back(pos[k]);
                Vec val = evalExpr(x, cx, sub);
                for (size_t j = 0; j < where.size(); j++) {
                    if (val.isNull(j)) continue;
                    r.null[where[j]] = 0;
                    copyAt(r, where[j], val, j);
                }
            };
            const size_t nwhen = (e.args.size() - 1 - (e.hasElse ? 1 : 0)) / 2;
            for (size_t w = 0; w < nwhen && !open.empty(); w++) {
                std::vector<int32_t> sub;
                for (size_t k : open) sub.push_back(pos[k]);
                Vec o = evalExpr(*e.args[1 + 2 * w], cx, sub);
                Vec vs = scatterNew(v.type, open.size());
                for (size_t j = 0; j < open.size(); j++) copyAt(vs, j, v, open[j]);
                Vec eq = compareOp(Op::EQ, o, vs);
                std::vector<size_t> hit, rest;
                for (size_t j = 0; j < open.size(); j++) ((!eq.isNull(j) && eq.i[j]) ? hit : rest).push_back(open[j]);
                place(*e.args[2 + 2 * w], hit);
                open.swap(rest);
            }
            toElse.insert(toElse.end(), open.begin(), open.end());
            if (e.hasElse) place(*e.args.back(), toElse);
            bool any = false;
            for (auto x : r.null) any |= x != 0;
            if (!any) r.null.clear();
            return r;
        }
        case ExprKind::CALL: {
            // BytecodeUtils.generateFullInvocation (main/sql/gen/BytecodeUtils.java:244-310): arguments are evaluated left to right and
            // the call is left as soon as one is NULL — a later argument is not evaluated (and cannot raise) at that position
            std::vector<Vec> a;
            std::vector<char> dead(n, 0);
            bool anyDead = false;
            for (auto& x : e.args) {
                if (!anyDead) {
                    a.push_back(evalExpr(*x, cx, pos));
                } else {
                    std::vector<int32_t> live;
                    std::vector<size_t> at;
                    for (size_t k = 0; k < n; k++)
                        if (!dead[k]) { live.push_back(pos[k]); at.push_back(k); }
                    Vec full = scatterNew(x->type, n);
                    full.null.assign(n, 1);
                    if (!live.empty()) {
                        Vec part = evalExpr(*x, cx, live);
                        for (size_t j = 0; j < at.size(); j++) {
                            if (part.isNull(j)) continue;
                            full.null[at[j]] = 0;
                            copyAt(full, at[j], part, j);
                        }
                    }
                    a.push_back(std::move(full));
                }
                for (size_t k = 0; k < n; k++)
                    if (a.back().isNull(k)) { dead[k] = 1; anyDead = true; }
            }
            switch (e.op) {
                case Op::EQ: case Op::NE: case Op::LT: case Op::LE: case Op::GT: case Op::GE: return compareOp(e.op, a[0], a[1]);
                case Op::NOT: {
                    Vec r = a[0];
                    for (size_t k = 0; k < n; k++) r.i[k] = !r.i[k];
                    return r;
                }
                case Op::NEGATE: {
                    Vec r = a[0];
                    for (size_t k = 0; k < n; k++) {
                        if (r.isNull(k)) continue;
                        if (e.type.kind == TypeKind::DOUBLE) r.d[k] = -r.d[k];
                        else if (e.type.isLongDecimal()) { if (!isZero(r.q[k]) || true) r.q[k].hi ^= SIGN_LONG_MASK; }  // UnscaledDecimal128Arithmetic.negate
                        else if (e.type.kind == TypeKind::BIGINT) {
                            if (r.i[k] == INT64_MIN) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "bigint negation overflow");
                            r.i[k] = -r.i[k];
                        } else if (e.type.kind == TypeKind::INTEGER) {
                            if (r.i[k] == INT32_MIN) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "integer negation overflow");
                            r.i[k] = -r.i[k];
                        } else r.i[k] = -r.i[k];
                    }
                    return r;
                }
                case Op::CAST: return evalCast(e, a[0]);
                default: return evalArith(e, a[0], a[1]);
            }
        }
    }
    throw OracleError(PGPU_NOT_SUPPORTED, "expression");
}


// ------------------------------------------------------------------------------------------- GroupByHash
static inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }
static inline uint64_t longHash(int64_t v) {  // spi/type/AbstractLongType.java:126-130
    return rotl64((uint64_t)v * 0xC2B2AE3D27D4EB4Full, 31) * 0x9E3779B185EBCA87ull;
}
static inline uint64_t murmurHash3(uint64_t x) {  // it.unimi.dsi.fastutil.HashCommon.murmurHash3 (fastutil 8.3.0, not vendored)
    x ^= x >> 33; x *= 0xff51afd7ed558ccdull;
    x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull;
    x ^= x >> 33;
    return x;
}

static uint64_t valueHash(const Vec& v, size_t k) {
    if (v.isNull(k)) return 0;  // type/TypeUtils.java:42 NULL_HASH_CODE
    switch (v.type.kind) {
        case TypeKind::BOOLEAN: return v.i[k] ? 1231 : 1237;  // spi/type/BooleanType.java hash
        case TypeKind::DOUBLE: {  // spi/type/DoubleType.java:163-170
            double d = v.d[k];
            if (d == 0) d = 0;
            if (d != d) d = std::numeric_limits<double>::quiet_NaN();
            int64_t b;
            memcpy(&b, &d, 8);
            if (d != d) b = 0x7ff8000000000000ll;
            return longHash(b);
        }
        case TypeKind::VARCHAR:
        case TypeKind::VARBINARY: {  // reference: XxHash64 of the bytes (airlift slice, not vendored); placement only -> FNV-1a here
            uint64_t h = 0xcbf29ce484222325ull;
            for (unsigned char c : v.s[k]) { h ^= c; h *= 0x100000001b3ull; }
            return h;
        }
        case TypeKind::DECIMAL:
            if (v.type.isLongDecimal()) return longHash((int64_t)v.q[k].lo) * 31 + longHash((int64_t)getLong1(v.q[k]) ^ (isZero(v.q[k]) ? 0 : (int64_t)(v.q[k].hi >> 63)));
            return longHash(v.i[k]);
        default: return longHash(v.i[k]);
    }
}

// IS NOT DISTINCT FROM between position k of `a` and position j of `b`
static bool notDistinct(const Vec& a, size_t k, const Vec& b, size_t j) {
    bool an = a.isNull(k), bn = b.isNull(j);
    if (an || bn) return an == bn;
    if (a.type.kind == TypeKind::DOUBLE) {  // spi/type/DoubleType.java:182-195
        double x = a.d[k], y = b.d[j];
        if (x != x && y != y) return true;
        return x == y;
    }
    if (a.type.isLongDecimal()) return compare(a.q[k], b.q[j]) == 0;
    if (a.type.isVarWidth()) return a.s[k] == b.s[j];
    return a.i[k] == b.i[j];
}

static void appendValue(Vec& dst, const Vec& src, size_t k) {
    size_t n = dst.size();
    if (dst.type.kind == TypeKind::DOUBLE) dst.d.push_back(src.isNull(k) ? 0.0 : src.d[k]);
    else if (dst.type.isLongDecimal()) dst.q.push_back(src.isNull(k) ? Dec128() : src.q[k]);
    else if (dst.type.isVarWidth()) dst.s.push_back(src.isNull(k) ? std::string() : src.s[k]);
    else dst.i.push_back(src.isNull(k) ? 0 : src.i[k]);
    if (src.isNull(k)) {
        if (dst.null.empty()) dst.null.assign(n, 0);
        dst.null.push_back(1);
    } else if (!dst.null.empty()) {
        dst.null.push_back(0);
    }
}

// MultiChannelGroupByHash.java:281-342 / BigintGroupByHash.java:213-323: open addressing, linear probing,
// murmurHash3(rawHash) & mask, fill 0.75, x2 rehash, group ids in first-seen order.
struct GroupByHash {
    std::vector<Vec> keys;            // one Vec per key channel, indexed by group id
    std::vector<int32_t> table;       // group id or -1
    std::vector<uint64_t> rawHashByGroup;
    uint64_t mask = 0;
    int64_t maxFill = 0;
    int64_t groupCount = 0;

    void init(const std::vector<Type>& types, int64_t expected) {
        keys.clear();
        for (auto& t : types) {
            Vec v;
            v.type = t;
            keys.push_back(v);
        }
        uint64_t size = 2;
        uint64_t need = (uint64_t)std::ceil((double)std::max<int64_t>(expected, 1) / 0.75);  // HashCommon.arraySize(expected, 0.75f)
        while (size < need) size <<= 1;
        table.assign(size, -1);
        mask = size - 1;
        maxFill = (int64_t)(size * 0.75);
        groupCount = 0;
        rawHashByGroup.clear();
    }

    void rehash() {
        uint64_t size = table.size() * 2;
        if (size > (1ull << 30)) throw OracleError(PGPU_INSUFFICIENT_RESOURCES, "Size of hash table cannot exceed 1 billion entries");
        table.assign(size, -1);
        mask = size - 1;
        maxFill = (int64_t)(size * 0.75);
        for (int64_t g = 0; g < groupCount; g++) {
            uint64_t p = murmurHash3(rawHashByGroup[g]) & mask;
            while (table[p] != -1) p = (p + 1) & mask;
            table[p] = (int32_t)g;
        }
    }

    int64_t putIfAbsent(const std::vector<Vec>& row, size_t k) {
        uint64_t h = 0;  // InterpretedHashGenerator.java:66-74: result = result * 31 + hash(channel)
        for (auto& c : row) h = h * 31 + valueHash(c, k);
        uint64_t p = murmurHash3(h) & mask;
        for (;;) {
            int32_t g = table[p];
            if (g == -1) break;
            if (rawHashByGroup[g] == h) {
                bool eq = true;
                for (size_t c = 0; c < row.size() && eq; c++) eq = notDistinct(row[c], k, keys[c], (size_t)g);
                if (eq) return g;
            }
            p = (p + 1) & mask;
        }
        int64_t g = groupCount++;
        table[p] = (int32_t)g;
        rawHashByGroup.push_back(h);
        for (size_t c = 0; c < row.size(); c++) appendValue(keys[c], row[c], k);
        if (groupCount >= maxFill) rehash();
        return g;
    }
};


// ------------------------------------------------------------------------------------------- accumulators
struct AccState {
    int64_t count = 0;     // LongDoubleState.first / LongAndDoubleState.long / LongState / decimal row counter
    double dsum = 0.0;
    int64_t lsum = 0;
    Dec128 dec;            // LongDecimalWithOverflowState.longDecimal
    bool hasDec = false;   // longDecimal != null
    int64_t overflow = 0;
    bool isNull = true;    // Nullable{Long,Double}State
    int64_t lval = 0;
    double dval = 0.0;
    Dec128 qval;           // min / max over long DECIMAL (NullableInt128State-like: the value as the block holds it)
    std::string sval;      // min / max over VARCHAR (SliceState)
};

struct Accumulator {
    AggSpec spec;
    std::vector<AccState> st;

    void ensure(int64_t groups) {
        if ((int64_t)st.size() < groups) st.resize((size_t)groups);
    }

    // GroupedAccumulator.addInput loop (AccumulatorCompiler.java:490-569): position order, mask, null-skip
    void addInput(const int64_t* gid, size_t n, const Vec* arg, const Vec* mask) {
        const Type& t = spec.inType;
        for (size_t k = 0; k < n; k++) {
            if (mask && (mask->isNull(k) || !mask->i[k])) continue;
            if (arg && arg->isNull(k)) continue;
            AccState& s = st[(size_t)gid[k]];
            switch (spec.fn) {
                case AggFn::COUNT: s.count++; break;  // CountAggregation.java:38-41, CountColumn.java
                case AggFn::SUM:
                    if (t.kind == TypeKind::DOUBLE) {  // DoubleSumAggregation.java:34-38
                        s.count++;
                        s.dsum += arg->d[k];
                    } else if (t.kind == TypeKind::BIGINT) {  // LongSumAggregation.java:35-39 (BigintOperators.add throws)
                        s.count++;
                        s.lsum = addExact(s.lsum, arg->i[k], "addition");
                    } else {  // DecimalSumAggregation.java:141-161
                        s.hasDec = true;
                        s.count++;
                        s.overflow += addWithOverflow(s.dec, decOf(*arg, k), s.dec);
                    }
                    break;
                case AggFn::AVG:
                    if (t.kind == TypeKind::DECIMAL) {  // DecimalAverageAggregation.java:153-177
                        s.count++;
                        s.hasDec = true;
                        s.overflow += addWithOverflow(s.dec, decOf(*arg, k), s.dec);
                    } else {  // AverageAggregations.java:35-46: the sum is a double for both input types
                        s.count++;
                        s.dsum += t.kind == TypeKind::DOUBLE ? arg->d[k] : (double)arg->i[k];
                    }
                    break;
                case AggFn::MIN:
                case AggFn::MAX: {  // AbstractMinMaxAggregationFunction.java:274-306 + util/MinMaxCompare.java
                    const bool isMax = spec.fn == AggFn::MAX;
                    if (t.kind == TypeKind::DOUBLE) {
                        double v = arg->d[k];
                        if (s.isNull) { s.isNull = false; s.dval = v; break; }
                        bool take;
                        if (isMax) take = (v > s.dval) || (s.dval != s.dval);  // maxDouble
                        else {  // Double.compare(v, state) < 0
                            auto key = [](double x) -> int64_t {
                                if (x != x) return 0x7ff8000000000000ll;
                                int64_t b;
                                memcpy(&b, &x, 8);
                                return b ^ ((b >> 63) & 0x7fffffffffffffffll);
                            };
                            take = key(v) < key(s.dval);
                        }
                        if (take) s.dval = v;
                    } else if (t.kind == TypeKind::VARCHAR) {  // SliceState, compared with Slice.compareTo: unsigned bytes, a proper prefix first
                        const std::string& v = arg->s[k];
                        if (s.isNull) { s.isNull = false; s.sval = v; break; }
                        const size_t m = std::min(v.size(), s.sval.size());
                        int c = memcmp(v.data(), s.sval.data(), m);
                        if (c == 0) c = v.size() < s.sval.size() ? -1 : v.size() > s.sval.size() ? 1 : 0;
                        if (isMax ? c > 0 : c < 0) s.sval = v;
                    } else if (t.isLongDecimal()) {  // Int128 state compared with UnscaledDecimal128Arithmetic.compare (decimal128.h)
                        const Dec128& v = arg->q[k];
                        if (s.isNull) { s.isNull = false; s.qval = v; break; }
                        const int c = compare(v, s.qval);
                        if (isMax ? c > 0 : c < 0) s.qval = v;
                    } else {
                        int64_t v = arg->i[k];
                        if (s.isNull) { s.isNull = false; s.lval = v; break; }
                        if (isMax ? v > s.lval : v < s.lval) s.lval = v;
                    }
                    break;
                }
            }
        }
    }

    // addIntermediate = deserialize + combine (AccumulatorCompiler.java:674-767); `ch` = the intermediate channels
    void addIntermediate(const int64_t* gid, size_t n, const std::vector<const Vec*>& ch) {
        const Type& t = spec.inType;
        for (size_t k = 0; k < n; k++) {
            AccState& s = st[(size_t)gid[k]];
            switch (spec.fn) {
                case AggFn::COUNT:
                    if (!ch[0]->isNull(k)) s.count += ch[0]->i[k];
                    break;
                case AggFn::SUM:
                case AggFn::AVG:
                    if (t.kind == TypeKind::DECIMAL) {  // DecimalSumAggregation.combine :163-175, DecimalAverageAggregation.combine :179-194
                        if (ch[0]->isNull(k)) break;
                        const std::string& b = ch[0]->s[k];
                        const uint64_t* w = (const uint64_t*)b.data();
                        int o = spec.fn == AggFn::AVG ? 1 : 0;
                        if (spec.fn == AggFn::AVG) s.count += (int64_t)w[0];
                        else s.count += 1;
                        int64_t ov = (int64_t)w[o];
                        Dec128 other;
                        other.lo = w[o + 1];
                        other.hi = w[o + 2];
                        if (!s.hasDec) { s.hasDec = true; s.dec = other; }
                        else ov += addWithOverflow(s.dec, other, s.dec);
                        s.overflow += ov;
                    } else if (spec.fn == AggFn::AVG) {  // ROW(double, bigint)
                        if (!ch[0]->isNull(k)) s.dsum += ch[0]->d[k];
                        if (!ch[1]->isNull(k)) s.count += ch[1]->i[k];
                    } else {  // ROW(bigint count, sum)
                        if (!ch[0]->isNull(k)) s.count += ch[0]->i[k];
                        if (!ch[1]->isNull(k)) {
                            if (t.kind == TypeKind::DOUBLE) s.dsum += ch[1]->d[k];
                            else s.lsum = addExact(s.lsum, ch[1]->i[k], "addition");
                        }
                    }
                    break;
                case AggFn::MIN:
                case AggFn::MAX: {
                    if (ch[0]->isNull(k)) break;
                    Vec one = *ch[0];  // reuse addInput's compare on a single value
                    (void)one;
                    int64_t g = gid[k];
                    Vec v;
                    v.type = ch[0]->type;
                    if (t.kind == TypeKind::DOUBLE) v.d.push_back(ch[0]->d[k]);
                    else if (t.isLongDecimal()) v.q.push_back(ch[0]->q[k]);
                    else if (t.kind == TypeKind::VARCHAR) v.s.push_back(ch[0]->s[k]);
                    else v.i.push_back(ch[0]->i[k]);
                    addInput(&g, 1, &v, nullptr);
                    break;
                }
            }
        }
    }
};

// BigDecimal(sum).divide(count, scale, ROUND_HALF_UP) on the exact sum overflow * 2^127 + value
// (DecimalAverageAggregation.average :218-227). Exact long division on a 192-bit magnitude.
static i128 decimalAverage(const AccState& s) {
    // signed 192-bit value = overflow * 2^127 + signed(dec), as (neg, mag[3] little-endian 64-bit limbs)
    auto add192 = [](uint64_t (&a)[3], const uint64_t (&b)[3]) {
        unsigned __int128 c = 0;
        for (int i = 0; i < 3; i++) { c += (unsigned __int128)a[i] + b[i]; a[i] = (uint64_t)c; c >>= 64; }
    };
    auto neg192 = [](uint64_t (&a)[3]) {
        unsigned __int128 c = 1;
        for (int i = 0; i < 3; i++) { c += (unsigned __int128)(~a[i]); a[i] = (uint64_t)c; c >>= 64; }
    };
    uint64_t v[3] = {s.dec.lo, getLong1(s.dec), 0};
    if (isNegative(s.dec)) neg192(v);
    // overflow * 2^127 in two's complement 192
    i128 ovs = (i128)s.overflow << 63;  // overflow * 2^63, placed at limb 1 => * 2^127
    uint64_t o[3] = {0, (uint64_t)(u128)ovs, (uint64_t)((u128)ovs >> 64)};
    add192(v, o);
    bool neg = (v[2] >> 63) != 0;
    if (neg) neg192(v);
    // divide magnitude by count with remainder
    uint64_t cnt = (uint64_t)s.count;
    uint64_t qd[3];
    unsigned __int128 rem = 0;
    for (int i = 2; i >= 0; i--) {
        unsigned __int128 cur = (rem << 64) | v[i];
        qd[i] = (uint64_t)(cur / cnt);
        rem = cur % cnt;
    }
    if ((unsigned __int128)rem * 2 >= cnt) {  // HALF_UP
        uint64_t one[3] = {1, 0, 0};
        add192(qd, one);
    }
    if (qd[2] != 0 || (qd[1] >> 63)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
    i128 r = (i128)(((u128)qd[1] << 64) | qd[0]);
    return neg ? -r : r;
}


// ------------------------------------------------------------------------------------------- the operator
static const int kBatch = 8192;  // PageProcessor.MAX_BATCH_SIZE (operator/project/PageProcessor.java:56)

struct OutBuffers {  // library-owned output page
    std::vector<Vec> cols;
    std::vector<std::vector<uint8_t>> values, nulls;
    std::vector<std::vector<int32_t>> offsets;
    int64_t rows = 0;
};

struct Oracle {
    std::unique_ptr<Plan> plan;
    std::vector<Type> projTypes;
    GroupByHash hash;
    std::vector<Accumulator> accs;
    bool finishing = false, finished = false, sawInput = false;
    std::vector<Vec> pending;      // filter + project: projected rows of the page just added
    bool pendingReady = false;     // ... waiting for the look getOutput takes (the GPU operator learns the row count there)
    // MergePages (operator/project/MergePages.java:87-164): `builder` is its PageBuilder, `ready` the pages it has let go
    std::deque<std::vector<Vec>> ready;
    std::vector<Vec> builder;
    int64_t frontServed = 0;       // rows of ready.front() already handed out (max-output-rows)
    std::vector<Vec> result;       // aggregation result rows not yet served
    bool resultReady = false;
    int64_t resultServed = 0;
    OutBuffers out;
    std::string err;
    int64_t groupCountGlobal = 1;

    bool hasGroups() const { return !plan->groupChannels.empty(); }

    void resetAgg() {
        std::vector<Type> kt;
        for (int c : plan->groupChannels) kt.push_back(projTypes[c]);
        hash.init(kt, plan->expectedGroups);
        accs.clear();
        for (auto& a : plan->aggs) {
            Accumulator acc;
            acc.spec = a;
            accs.push_back(acc);
        }
        if (!hasGroups())
            for (auto& a : accs) a.ensure(1);
    }

    void init() {
        projTypes = plan->projectedTypes();
        if (plan->hasAgg) resetAgg();
    }

    // one batch: filter -> selected positions -> projections -> aggregation or output rows
    void processBatch(const std::vector<BlockView>& cols, int64_t base, int n) {
        EvalCtx cx{&cols, base};
        std::vector<int32_t> all((size_t)n);
        for (int k = 0; k < n; k++) all[k] = k;
        std::vector<int32_t> selected;
        if (plan->filter) {
            // generated filter loop (PageFunctionCompiler.java:485-497): selected[position] = filter(...) with NULL => false (:542),
            // then PageFilter.positionsArrayToSelectedPositions (PageFilter.java:27-50): ascending position list
            Vec f = evalExpr(*plan->filter, cx, all);
            for (int k = 0; k < n; k++)
                if (!f.isNull(k) && f.i[k]) selected.push_back(k);
        } else {
            selected.swap(all);
        }
        if (selected.empty()) return;
        size_t m = selected.size();
        // projections at the selected positions only (PageFunctionCompiler.java:283-319)
        std::vector<Vec> proj(plan->projections.size());
        std::vector<bool> have(plan->projections.size(), false);
        auto need = [&](int c) {
            if (!have[c]) {
                proj[c] = evalExpr(*plan->projections[c], cx, selected);
                have[c] = true;
            }
        };
        if (!plan->hasAgg) {
            for (size_t c = 0; c < proj.size(); c++) need((int)c);
            if (pending.empty()) {
                pending.resize(proj.size());
                for (size_t c = 0; c < proj.size(); c++) pending[c].type = projTypes[c];
            }
            for (size_t c = 0; c < proj.size(); c++)
                for (size_t k = 0; k < m; k++) appendValue(pending[c], proj[c], k);
            return;
        }
        // group ids (GroupByHash.getGroupIds), then one pass per aggregate (InMemoryHashAggregationBuilder.processPage :139-155)
        std::vector<int64_t> gid(m, 0);
        if (hasGroups()) {
            std::vector<Vec> keyRow;
            for (int c : plan->groupChannels) {
                need(c);
                keyRow.push_back(proj[c]);
            }
            for (size_t k = 0; k < m; k++) gid[k] = hash.putIfAbsent(keyRow, k);
            for (auto& a : accs) a.ensure(hash.groupCount);
        }
        const bool fin = inputIsIntermediate(plan->step);
        for (auto& a : accs) {
            if (fin) {
                size_t ni = Plan::intermediateTypes(a.spec).size();
                std::vector<const Vec*> ch;
                for (size_t j = 0; j < ni; j++) {
                    need(a.spec.arg + (int)j);
                    ch.push_back(&proj[a.spec.arg + j]);
                }
                a.addIntermediate(gid.data(), m, ch);
            } else {
                const Vec *arg = nullptr, *mask = nullptr;
                if (a.spec.hasArg) { need(a.spec.arg); arg = &proj[a.spec.arg]; }
                if (a.spec.mask >= 0) { need(a.spec.mask); mask = &proj[a.spec.mask]; }
                a.addInput(gid.data(), m, arg, mask);
            }
        }
    }

    void addPage(const pgpu_column* cols, int ncols, int64_t rows) {
        if (ncols != (int)plan->input.size()) throw OracleError(PGPU_INVALID_ARGUMENT, "channel count");
        sawInput = true;
        std::vector<BlockView> views;
        for (int c = 0; c < ncols; c++) views.push_back(viewOf(cols[c], plan->input[c]));
        for (int64_t base = 0; base < rows; base += kBatch) processBatch(views, base, (int)std::min<int64_t>(kBatch, rows - base));
        if (!plan->hasAgg) pendingReady = true;
    }

    // InMemoryHashAggregationBuilder.buildResult (:272-298) / AggregationOperator.getOutput (:164-186): group ids in
    // first-seen order; aggregator.evaluate(groupId, output) = final value or serialized intermediate state
    void buildResult() {
        std::vector<Type> ot = plan->outputTypes();
        result.assign(ot.size(), Vec());
        for (size_t c = 0; c < ot.size(); c++) result[c].type = ot[c];
        int64_t groups = hasGroups() ? hash.groupCount : 1;
        const bool partial = outputIsIntermediate(plan->step);
        for (int64_t g = 0; g < groups; g++) {
            size_t oc = 0;
            for (size_t c = 0; c < plan->groupChannels.size(); c++) appendValue(result[oc++], hash.keys[c], (size_t)g);
            for (auto& a : accs) {
                a.ensure(groups);
                const AccState& s = a.st[(size_t)g];
                const Type& t = a.spec.inType;
                auto putI = [&](Vec& v, int64_t x, bool isNull) { Vec one; one.type = v.type; one.i.push_back(x); if (isNull) one.null.push_back(1); appendValue(v, one, 0); };
                auto putD = [&](Vec& v, double x, bool isNull) { Vec one; one.type = v.type; one.d.push_back(x); if (isNull) one.null.push_back(1); appendValue(v, one, 0); };
                auto putQ = [&](Vec& v, Dec128 x, bool isNull) { Vec one; one.type = v.type; one.q.push_back(x); if (isNull) one.null.push_back(1); appendValue(v, one, 0); };
                auto putS = [&](Vec& v, const std::string& x, bool isNull) { Vec one; one.type = v.type; one.s.push_back(x); if (isNull) one.null.push_back(1); appendValue(v, one, 0); };
                switch (a.spec.fn) {
                    case AggFn::COUNT: putI(result[oc++], s.count, false); break;
                    case AggFn::SUM:
                    case AggFn::AVG:
                        if (t.kind == TypeKind::DECIMAL) {
                            if (partial) {  // LongDecimalWithOverflow[AndLong]StateSerializer.serialize
                                std::string b;
                                if (a.spec.fn == AggFn::AVG) b.append((const char*)&s.count, 8);
                                b.append((const char*)&s.overflow, 8);
                                b.append((const char*)&s.dec.lo, 8);
                                b.append((const char*)&s.dec.hi, 8);
                                putS(result[oc++], b, !s.hasDec);
                            } else if (a.spec.fn == AggFn::SUM) {  // DecimalSumAggregation.outputLongDecimal :177-189
                                if (s.hasDec) {
                                    if (s.overflow != 0) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
                                    if (exceedsOrEqualTenToThirtyEight(s.dec)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
                                }
                                putQ(result[oc++], s.dec, !s.hasDec);
                            } else {  // DecimalAverageAggregation.output{Short,Long}Decimal :196-216
                                if (s.count == 0) {
                                    if (a.spec.outType.isShortDecimal()) putI(result[oc++], 0, true);
                                    else putQ(result[oc++], Dec128(), true);
                                } else {
                                    i128 avg = decimalAverage(s);
                                    if (a.spec.outType.isShortDecimal()) {
                                        if (avg > (i128)INT64_MAX || avg < (i128)INT64_MIN) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
                                        putI(result[oc++], (int64_t)avg, false);
                                    } else {
                                        Dec128 q = fromI128(avg);
                                        if (exceedsOrEqualTenToThirtyEight(q)) throw OracleError(PGPU_NUMERIC_VALUE_OUT_OF_RANGE, "Decimal overflow");
                                        putQ(result[oc++], q, false);
                                    }
                                }
                            }
                        } else if (a.spec.fn == AggFn::AVG) {
                            if (partial) { putD(result[oc++], s.dsum, false); putI(result[oc++], s.count, false); }
                            else putD(result[oc++], s.count ? s.dsum / (double)s.count : 0.0, s.count == 0);  // AverageAggregations.output :66-79
                        } else if (t.kind == TypeKind::DOUBLE) {
                            if (partial) { putI(result[oc++], s.count, false); putD(result[oc++], s.dsum, false); }
                            else putD(result[oc++], s.dsum, s.count == 0);  // DoubleSumAggregation.output :54-62
                        } else {
                            if (partial) { putI(result[oc++], s.count, false); putI(result[oc++], s.lsum, false); }
                            else putI(result[oc++], s.lsum, s.count == 0);  // LongSumAggregation.output :54-63
                        }
                        break;
                    case AggFn::MIN:
                    case AggFn::MAX:
                        if (t.kind == TypeKind::DOUBLE) putD(result[oc++], s.dval, s.isNull);
                        else if (t.isLongDecimal()) putQ(result[oc++], s.qval, s.isNull);
                        else if (t.kind == TypeKind::VARCHAR) putS(result[oc++], s.sval, s.isNull);
                        else putI(result[oc++], s.lval, s.isNull);
                        break;
                }
            }
        }
        resultReady = true;
        resultServed = 0;
    }

    static int64_t pageRows(const std::vector<Vec>& pg) { return pg.empty() ? 0 : (int64_t)pg[0].size(); }

    // Page.getSizeInBytes = sum of Block.getSizeInBytes: fixed width (width + 1 null byte) per position
    // (LongArrayBlock.java getSizeInBytes), var width the bytes + (4 offset + 1 null) per position (VariableWidthBlock.java)
    static int64_t pageBytes(const std::vector<Vec>& pg) {
        int64_t b = 0;
        for (auto& v : pg) {
            int64_t n = (int64_t)v.size();
            if (v.type.isVarWidth()) {
                b += n * 5;
                for (auto& x : v.s) b += (int64_t)x.size();
            } else {
                b += n * (v.type.width() + 1);
            }
        }
        return b;
    }

    void flushBuilder() {
        ready.push_back(std::move(builder));
        builder.clear();
    }

    // MergePages.java:112-151 on the page addPage left in `pending`
    void mergePagesProcess() {
        std::vector<Vec> pg;
        pg.swap(pending);
        if (pg.empty()) {
            pg.resize(projTypes.size());
            for (size_t c = 0; c < pg.size(); c++) pg[c].type = projTypes[c];
        }
        const bool merging = plan->mergeMinRows > 0 || plan->mergeMinBytes > 0;
        if (!merging || pageRows(pg) >= plan->mergeMinRows || pageBytes(pg) >= plan->mergeMinBytes) {
            if (pageRows(builder) > 0) flushBuilder();  // the buffered rows first, then the page as it is (:135-142)
            ready.push_back(std::move(pg));
            return;
        }
        if (builder.empty()) {
            builder.resize(pg.size());
            for (size_t c = 0; c < pg.size(); c++) builder[c].type = pg[c].type;
        }
        for (size_t c = 0; c < pg.size(); c++)          // appendPage (:153-162)
            for (size_t k = 0; k < pg[c].size(); k++) appendValue(builder[c], pg[c], k);
        if (pageBytes(builder) >= plan->mergeMaxBytes) flushBuilder();  // PageBuilder.isFull (:147-149)
    }

    bool partialFlushDue() const {
        return plan->hasAgg && outputIsIntermediate(plan->step) && plan->maxPartialGroups > 0 && hasGroups() && hash.groupCount >= plan->maxPartialGroups;
    }

    bool needsInput() const {  // HashAggregationOperator.needsInput :367-378
        if (finishing || finished) return false;
        if (pendingReady || !ready.empty() || resultReady) return false;
        if (partialFlushDue()) return false;
        return true;
    }

    // returns false for "no page now" (Java null)
    bool getOutput(std::vector<Vec>& page, int64_t& rows) {
        if (finished) return false;
        if (!plan->hasAgg) {
            if (pendingReady) {
                pendingReady = false;
                mergePagesProcess();
            }
            if (ready.empty()) {
                if (!finishing) return false;
                if (pageRows(builder) == 0) {  // MergePages.java:125-129
                    finished = true;
                    return false;
                }
                flushBuilder();                // MergePages.java:131
            }
            const std::vector<Vec>& front = ready.front();
            int64_t n = pageRows(front) - frontServed;
            if (plan->maxOutputRows > 0) n = std::min<int64_t>(n, plan->maxOutputRows);
            page.assign(front.size(), Vec());
            for (size_t c = 0; c < front.size(); c++) {
                page[c].type = front[c].type;
                for (int64_t k = 0; k < n; k++) appendValue(page[c], front[c], (size_t)(frontServed + k));
            }
            rows = n;
            frontServed += n;
            if (frontServed >= pageRows(front)) {
                ready.pop_front();
                frontServed = 0;
            }
            return true;
        }
        if (!resultReady) {
            if (!finishing && !partialFlushDue()) return false;
            if (hasGroups() && !sawInput) {  // HashAggregationOperator.getOutput :484-489
                finished = true;
                return false;
            }
            buildResult();
        }
        int64_t total = result.empty() ? 0 : (int64_t)result[0].size();
        if (!hasGroups() && result.empty()) total = 0;
        int64_t remaining = total - resultServed;
        if (remaining <= 0) {
            resultReady = false;
            if (finishing) finished = true;
            else resetAgg();
            return false;
        }
        int64_t n = remaining;
        if (plan->maxOutputRows > 0) n = std::min<int64_t>(n, plan->maxOutputRows);
        page.assign(result.size(), Vec());
        for (size_t c = 0; c < result.size(); c++) {
            page[c].type = result[c].type;
            for (int64_t k = 0; k < n; k++) appendValue(page[c], result[c], (size_t)(resultServed + k));
        }
        rows = n;
        resultServed += n;
        if (resultServed >= total) {
            resultReady = false;
            if (finishing) finished = true;
            else resetAgg();
        }
        return true;
    }

    void fillOut(const std::vector<Vec>& page, int64_t rows, pgpu_column* cols) {
        size_t nc = page.size();
        out.values.assign(nc, {});
        out.nulls.assign(nc, {});
        out.offsets.assign(nc, {});
        out.rows = rows;
        for (size_t c = 0; c < nc; c++) {
            const Vec& v = page[c];
            pgpu_column& o = cols[c];
            memset(&o, 0, sizeof o);
            o.encoding = PGPU_ENC_PLAIN;
            o.type_kind = (int32_t)v.type.kind;
            o.precision = v.type.precision;
            o.scale = v.type.scale;
            o.position_count = rows;
            out.nulls[c].assign((size_t)std::max<int64_t>(rows, 1), 0);
            for (int64_t k = 0; k < rows; k++) out.nulls[c][k] = v.isNull((size_t)k);
            auto& buf = out.values[c];
            if (v.type.isVarWidth()) {
                out.offsets[c].assign((size_t)rows + 1, 0);
                for (int64_t k = 0; k < rows; k++) {
                    if (!v.isNull((size_t)k)) buf.insert(buf.end(), v.s[k].begin(), v.s[k].end());
                    out.offsets[c][k + 1] = (int32_t)buf.size();
                }
                if (buf.empty()) buf.push_back(0);
                o.offsets = out.offsets[c].data();
            } else {
                int w = v.type.width();
                buf.assign((size_t)std::max<int64_t>(rows, 1) * w, 0);
                for (int64_t k = 0; k < rows; k++) {
                    unsigned char* dst = buf.data() + (size_t)k * w;
                    if (v.type.kind == TypeKind::DOUBLE) memcpy(dst, &v.d[k], 8);
                    else if (v.type.isLongDecimal()) { memcpy(dst, &v.q[k].lo, 8); memcpy(dst + 8, &v.q[k].hi, 8); }
                    else if (w == 8) memcpy(dst, &v.i[k], 8);
                    else if (w == 4) { int32_t x = (int32_t)v.i[k]; memcpy(dst, &x, 4); }
                    else { int8_t x = (int8_t)v.i[k]; memcpy(dst, &x, 1); }
                }
            }
            o.values = buf.data();
            o.nulls = out.nulls[c].data();
        }
    }

};


}  // namespace oracle

// ================================================================================================ C API (ctypes)
using namespace oracle;

#define ORACLE_GUARD(o, body)                                      \
    try {                                                          \
        body;                                                      \
        return PGPU_OK;                                            \
    } catch (const OracleError& e) {                               \
        (o)->err = e.what();                                       \
        return e.status;                                           \
    } catch (const PlanError& e) {                                 \
        (o)->err = e.what();                                       \
        return e.status;                                           \
    } catch (const ArithmeticOverflow& e) {                        \
        (o)->err = e.what();                                       \
        return PGPU_NUMERIC_VALUE_OUT_OF_RANGE;                    \
    }

extern "C" {

int oracle_create(const char* plan, size_t len, void** out, char* err, size_t errcap) {
    *out = nullptr;
    try {
        auto o = std::make_unique<Oracle>();
        o->plan = readPlan(std::string(plan, len));
        o->init();
        *out = o.release();
        return PGPU_OK;
    } catch (const PlanError& e) {
        if (err && errcap) snprintf(err, errcap, "%s", e.what());
        return e.status;
    }
}
void oracle_destroy(void* h) { delete (Oracle*)h; }
const char* oracle_last_error(void* h) { return ((Oracle*)h)->err.c_str(); }
int oracle_needs_input(void* h) { return ((Oracle*)h)->needsInput() ? 1 : 0; }
int oracle_is_finished(void* h) { return ((Oracle*)h)->finished ? 1 : 0; }
int oracle_finish(void* h) {
    ((Oracle*)h)->finishing = true;
    return PGPU_OK;
}
int oracle_add_page(void* h, const pgpu_column* cols, int ncols, int64_t rows) {
    Oracle* o = (Oracle*)h;
    if (o->finishing) { o->err = "Operator is already finishing"; return PGPU_INVALID_STATE; }
    if (!o->needsInput()) { o->err = "Operator does not need input"; return PGPU_INVALID_STATE; }
    ORACLE_GUARD(o, o->addPage(cols, ncols, rows));
}
int oracle_get_output(void* h, pgpu_column* cols, int max_cols, int* ncols, int64_t* rows) {
    Oracle* o = (Oracle*)h;
    *rows = -1;
    *ncols = 0;
    ORACLE_GUARD(o, {
        std::vector<Vec> page;
        int64_t n = 0;
        if (o->getOutput(page, n)) {
            if ((int)page.size() > max_cols) throw OracleError(PGPU_INVALID_ARGUMENT, "out_cols too small");
            o->fillOut(page, n, cols);
            *ncols = (int)page.size();
            *rows = n;
        }
    });
}

// ---- GroupByHash alone (TestGroupByHash): group ids of a page of key channels, first-seen order
int oracle_group_ids(const char* types_plan, size_t len, const pgpu_column* cols, int ncols, int64_t rows, int64_t expected, int64_t* ids_out, int64_t* group_count) {
    try {
        // types_plan: "(plan (input t...))" — only the input types are used
        auto plan = readPlan(std::string(types_plan, len));
        GroupByHash gh;
        gh.init(plan->input, expected);
        std::vector<BlockView> views;
        for (int c = 0; c < ncols; c++) views.push_back(viewOf(cols[c], plan->input[c]));
        for (int64_t base = 0; base < rows; base += kBatch) {
            int n = (int)std::min<int64_t>(kBatch, rows - base);
            std::vector<int32_t> pos((size_t)n);
            for (int k = 0; k < n; k++) pos[k] = k;
            std::vector<Vec> row;
            for (auto& v : views) row.push_back(gather(v, pos.data(), (size_t)n, base));
            for (int k = 0; k < n; k++) ids_out[base + k] = gh.putIfAbsent(row, (size_t)k);
        }
        *group_count = gh.groupCount;
        return PGPU_OK;
    } catch (const std::exception&) {
        return PGPU_INVALID_ARGUMENT;
    }
}

// ---- UnscaledDecimal128Arithmetic known-answer hooks: operands / results are 2 x uint64 block layout
int oracle_dec_add_with_overflow(const uint64_t* a, const uint64_t* b, uint64_t* r, int64_t* overflow) {
    Dec128 x{a[0], a[1]}, y{b[0], b[1]}, z;
    *overflow = addWithOverflow(x, y, z);
    r[0] = z.lo;
    r[1] = z.hi;
    return 0;
}
int oracle_dec_binary(int op, const uint64_t* a, const uint64_t* b, uint64_t* r) {  // 0 add 1 subtract 2 multiply; 1 = overflow
    try {
        Dec128 x{a[0], a[1]}, y{b[0], b[1]};
        Dec128 z = op == 0 ? add(x, y) : op == 1 ? subtract(x, y) : multiply(x, y);
        r[0] = z.lo;
        r[1] = z.hi;
        return 0;
    } catch (const ArithmeticOverflow&) {
        return 1;
    }
}
int oracle_dec_rescale(const uint64_t* a, int factor, uint64_t* r) {
    try {
        Dec128 z = rescale(Dec128{a[0], a[1]}, factor);
        r[0] = z.lo;
        r[1] = z.hi;
        return 0;
    } catch (const ArithmeticOverflow&) {
        return 1;
    }
}
int oracle_dec_compare(const uint64_t* a, const uint64_t* b) { return compare(Dec128{a[0], a[1]}, Dec128{b[0], b[1]}); }
int oracle_dec_overflows(const uint64_t* a) { return exceedsOrEqualTenToThirtyEight(Dec128{a[0], a[1]}) ? 1 : 0; }

}  // extern "C"

namespace oracle {







}  // namespace oracle
