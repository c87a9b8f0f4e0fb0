// oracle/oracle_plan.h — TEST INFRASTRUCTURE: the oracle's OWN reader of the plan text the tests hand to both sides.
//
// Nothing here is shared with presto_b200/csrc/plan.cpp (the product's front end): the text is read by a second, independent
// implementation, type rules are restated from the reference, and NULL_IF / SWITCH stay what they are in the reference's IR
// (sql/relational/SpecialForm.java Form.NULL_IF / SWITCH / WHEN) and are EVALUATED as such (oracle.cpp evalExpr) instead of
// being lowered to IF chains the way the product does. A lowering bug in the product therefore shows up as a parity failure.
//
// Plan text (tests build it with presto_b200/operators.py plan_text and relational.py sexpr):
//   (plan (input T...) [(filter E)] [(project E...)] [(aggregate (step S) [(group c...)] [(expected-groups n)] (agg ...)...)]
//         [(options (k v)...)])
//   E := (in c T) | (const T lit) | (call op T E...) | (and E E...) | (or E E...) | (between E E E) | (in-list E C...)
//      | (is-null E) | (if T E E E) | (coalesce T E E...) | (null-if E E) | (switch T E (when E E)... [E])
// The oracle does not judge what the GPU supports: it reads every well-formed plan of this grammar.
#pragma once
#include <cctype>
#include <cerrno>
#include <climits>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/prestogpu.h"  // pgpu_status codes, pgpu_type_kind numbering

namespace oracle {

typedef __int128 i128;
typedef unsigned __int128 u128;

// numbering = pgpu_type_kind of include/prestogpu.h (the column descriptors both sides receive)
enum class TypeKind : int { BOOLEAN = 0, BIGINT = 1, INTEGER = 2, DATE = 3, DOUBLE = 4, DECIMAL = 5, VARCHAR = 6, VARBINARY = 7 };

struct Type {
    TypeKind kind = TypeKind::BIGINT;
    int precision = 0;  // DECIMAL precision / VARCHAR bound (0 = unbounded)
    int scale = 0;
    bool operator==(const Type& o) const { return kind == o.kind && precision == o.precision && scale == o.scale; }
    bool operator!=(const Type& o) const { return !(*this == o); }
    bool isDecimal() const { return kind == TypeKind::DECIMAL; }
    bool isShortDecimal() const { return isDecimal() && precision <= 18; }  // spi/type/Decimals.java MAX_SHORT_PRECISION = 18
    bool isLongDecimal() const { return isDecimal() && precision > 18; }
    bool isVarWidth() const { return kind == TypeKind::VARCHAR || kind == TypeKind::VARBINARY; }
    int width() const {  // bytes per position of the type's flat block (ByteArrayBlock / IntArrayBlock / LongArrayBlock / Int128ArrayBlock)
        switch (kind) {
            case TypeKind::BOOLEAN: return 1;
            case TypeKind::INTEGER:
            case TypeKind::DATE: return 4;
            case TypeKind::BIGINT:
            case TypeKind::DOUBLE: return 8;
            case TypeKind::DECIMAL: return precision <= 18 ? 8 : 16;
            default: return 0;
        }
    }
    std::string str() const {
        switch (kind) {
            case TypeKind::BOOLEAN: return "boolean";
            case TypeKind::BIGINT: return "bigint";
            case TypeKind::INTEGER: return "integer";
            case TypeKind::DATE: return "date";
            case TypeKind::DOUBLE: return "double";
            case TypeKind::DECIMAL: return "decimal(" + std::to_string(precision) + "," + std::to_string(scale) + ")";
            case TypeKind::VARCHAR: return precision ? "varchar(" + std::to_string(precision) + ")" : "varchar";
            default: return "varbinary";
        }
    }
    static Type boolean() { return Type{TypeKind::BOOLEAN, 0, 0}; }
    static Type bigint() { return Type{TypeKind::BIGINT, 0, 0}; }
    static Type integer() { return Type{TypeKind::INTEGER, 0, 0}; }
    static Type date() { return Type{TypeKind::DATE, 0, 0}; }
    static Type dbl() { return Type{TypeKind::DOUBLE, 0, 0}; }
    static Type decimal(int p, int s) { return Type{TypeKind::DECIMAL, p, s}; }
    static Type varchar(int n) { return Type{TypeKind::VARCHAR, n, 0}; }
    static Type varbinary() { return Type{TypeKind::VARBINARY, 0, 0}; }
};

enum class ExprKind { INPUT, CONST, CALL, AND, OR, BETWEEN, IN, IS_NULL, IF, COALESCE, NULL_IF, SWITCH };
enum class Op { ADD, SUBTRACT, MULTIPLY, DIVIDE, MODULUS, NEGATE, EQ, NE, LT, LE, GT, GE, NOT, CAST };

struct Expr {
    ExprKind kind = ExprKind::CONST;
    Type type;
    int channel = -1;      // INPUT
    bool isNull = false;   // CONST NULL
    int64_t i64 = 0;       // CONST: boolean / integer / date / bigint / short decimal value, or the bits of a DOUBLE
    i128 dec = 0;          // CONST: long decimal, two's complement
    Op op = Op::ADD;       // CALL
    // SWITCH: args = value, then (operand, result) pairs, then the else expression when hasElse
    bool hasElse = false;
    std::vector<std::unique_ptr<Expr>> args;
};

enum class AggFn { SUM, COUNT, AVG, MIN, MAX };
enum class Step { SINGLE, PARTIAL, FINAL, INTERMEDIATE };  // sql/planner/plan/AggregationNode.java:348-380
inline bool inputIsIntermediate(Step s) { return s == Step::FINAL || s == Step::INTERMEDIATE; }    // !Step.isInputRaw()
inline bool outputIsIntermediate(Step s) { return s == Step::PARTIAL || s == Step::INTERMEDIATE; } // Step.isOutputPartial()

struct AggSpec {
    AggFn fn = AggFn::COUNT;
    Type outType, inType;
    bool hasArg = false;  // false: count(*)
    int arg = -1;         // raw input: the argument's channel; intermediate input: the first state channel
    int mask = -1;
};

struct PlanError : std::runtime_error {
    int status;
    PlanError(int st, const std::string& m) : std::runtime_error(m), status(st) {}
};

struct Plan {
    std::vector<Type> input;
    std::unique_ptr<Expr> filter;
    std::vector<std::unique_ptr<Expr>> projections;  // identity when the text has no (project ...)
    bool hasAgg = false;
    Step step = Step::SINGLE;
    std::vector<int> groupChannels;
    int64_t expectedGroups = 10000;
    std::vector<AggSpec> aggs;
    int64_t maxPartialGroups = 0, maxOutputRows = 0;
    int64_t mergeMinRows = 0, mergeMinBytes = 0, mergeMaxBytes = 1 << 20;  // MergePages.java:60-84; PageBuilder.DEFAULT_MAX_PAGE_SIZE_IN_BYTES

    std::vector<Type> projectedTypes() const {
        std::vector<Type> t;
        for (auto& p : projections) t.push_back(p->type);
        return t;
    }
    // The channels one aggregate's state travels in between PARTIAL and FINAL. The reference serialises a multi-field state as one
    // RowType channel (state/StateCompiler.java:211, fields ordered primitives first, then by name, :588-630); both sides of this
    // repository flatten the fields into consecutive channels in that order:
    //   count                      LongState                         -> bigint
    //   sum(bigint)                LongLongState (count, sum)        -> bigint, bigint        (LongSumAggregation.java)
    //   sum(double)                LongDoubleState (count, sum)      -> bigint, double        (DoubleSumAggregation.java)
    //   avg(bigint|double)         LongAndDoubleState (double, long) -> double, bigint        (AverageAggregations.java)
    //   sum|avg(decimal)           LongDecimalWithOverflow[AndLong]State -> varbinary         (its serializers, :43-52)
    //   min|max(T)                 Nullable*State                    -> T
    static std::vector<Type> intermediateTypes(const AggSpec& a) {
        const bool dec = a.inType.isDecimal();
        if (a.fn == AggFn::COUNT) return {Type::bigint()};
        if (a.fn == AggFn::MIN || a.fn == AggFn::MAX) return {a.inType};
        if (dec) return {Type::varbinary()};
        if (a.fn == AggFn::AVG) return {Type::dbl(), Type::bigint()};
        return {Type::bigint(), a.inType.kind == TypeKind::DOUBLE ? Type::dbl() : Type::bigint()};
    }
    std::vector<Type> outputTypes() const {
        std::vector<Type> proj = projectedTypes(), out;
        if (!hasAgg) return proj;
        for (int g : groupChannels) out.push_back(proj[(size_t)g]);
        for (auto& a : aggs) {
            if (!outputIsIntermediate(step)) {
                out.push_back(a.outType);
                continue;
            }
            for (auto& t : intermediateTypes(a)) out.push_back(t);
        }
        return out;
    }
};

inline i128 pow10_i128(int n) {
    i128 v = 1;
    while (n-- > 0) v *= 10;
    return v;
}

std::unique_ptr<Plan> readPlan(const std::string& text);

// ================================================================================================ reader
namespace plan_reader {

constexpr int kInvalidArgument = PGPU_INVALID_ARGUMENT;
constexpr int kNotSupported = PGPU_NOT_SUPPORTED;

[[noreturn]] inline void malformed(const std::string& why) { throw PlanError(kInvalidArgument, "oracle: malformed plan: " + why); }

// A cursor over the text: lists are entered and left, atoms are taken one by one. No tree is built.
class Cursor {
    const std::string& s;
    size_t at = 0;
    void blanks() {
        for (;;) {
            while (at < s.size() && isspace((unsigned char)s[at])) at++;
            if (at < s.size() && s[at] == ';') {  // comment to the end of the line
                while (at < s.size() && s[at] != '\n') at++;
                continue;
            }
            return;
        }
    }

public:
    explicit Cursor(const std::string& text) : s(text) {}
    bool atOpen() { blanks(); return at < s.size() && s[at] == '('; }
    bool atClose() { blanks(); return at < s.size() && s[at] == ')'; }
    bool atEnd() { blanks(); return at >= s.size(); }
    void enter() {
        if (!atOpen()) malformed("'(' expected at offset " + std::to_string(at));
        at++;
    }
    void leave() {
        if (!atClose()) malformed("')' expected at offset " + std::to_string(at));
        at++;
    }
    std::string atom() {
        blanks();
        size_t b = at;
        while (at < s.size() && !isspace((unsigned char)s[at]) && s[at] != '(' && s[at] != ')') at++;
        if (at == b) malformed("a word expected at offset " + std::to_string(b));
        return s.substr(b, at - b);
    }
    // enters a list and returns its first word
    std::string enterNamed() {
        enter();
        return atom();
    }
    // the first word of the list ahead, without moving
    std::string peekName() {
        size_t keep = at;
        std::string n = atOpen() ? enterNamed() : std::string();
        at = keep;
        return n;
    }
    void skipList() {  // skip one balanced list
        enter();
        int depth = 1;
        while (depth > 0) {
            blanks();
            if (at >= s.size()) malformed("unbalanced parentheses");
            if (s[at] == '(') { depth++; at++; }
            else if (s[at] == ')') { depth--; at++; }
            else atom();
        }
    }
    int64_t integer(const char* what) {
        std::string a = atom();
        errno = 0;
        char* end = nullptr;
        long long v = strtoll(a.c_str(), &end, 0);
        if (errno || end == a.c_str() || *end) malformed(std::string("'") + a + "' is not an integer (" + what + ")");
        return v;
    }
};

inline Type readType(Cursor& c) {
    if (!c.atOpen()) {
        std::string w = c.atom();
        static const struct { const char* name; Type t; } simple[] = {
            {"boolean", Type::boolean()}, {"bigint", Type::bigint()}, {"integer", Type::integer()}, {"date", Type::date()},
            {"double", Type::dbl()}, {"varchar", Type::varchar(0)}, {"varbinary", Type::varbinary()}};
        for (auto& e : simple)
            if (w == e.name) return e.t;
        throw PlanError(kNotSupported, "oracle: type '" + w + "' is outside the restated path");
    }
    std::string name = c.enterNamed();
    Type t;
    if (name == "decimal") {
        int p = (int)c.integer("decimal precision"), s = (int)c.integer("decimal scale");
        if (p < 1 || p > 38 || s < 0 || s > p) malformed("decimal(" + std::to_string(p) + "," + std::to_string(s) + ")");  // spi/type/DecimalType.java:42-57
        t = Type::decimal(p, s);
    } else if (name == "varchar") {
        int n = (int)c.integer("varchar length");
        if (n < 0) malformed("varchar length");
        t = Type::varchar(n);
    } else {
        throw PlanError(kNotSupported, "oracle: type form '" + name + "' is outside the restated path");
    }
    c.leave();
    return t;
}

inline i128 readUnscaled(const std::string& lit) {
    size_t k = 0;
    bool minus = false;
    if (k < lit.size() && (lit[k] == '+' || lit[k] == '-')) minus = lit[k++] == '-';
    if (k == lit.size()) malformed("decimal literal '" + lit + "'");
    u128 mag = 0;
    for (; k < lit.size(); k++) {
        if (lit[k] < '0' || lit[k] > '9') malformed("decimal literal '" + lit + "'");
        if (mag > ((~(u128)0) >> 1) / 10) malformed("decimal literal out of range '" + lit + "'");
        mag = mag * 10 + (unsigned)(lit[k] - '0');
    }
    if (mag >> 127) malformed("decimal literal out of range '" + lit + "'");
    return minus ? -(i128)mag : (i128)mag;
}

}  // namespace plan_reader
}  // namespace oracle

#include "oracle_plan_expr.h"
