// plan.h — the plan language of the GPU pipeline: types, RowExpression subset, aggregates.
//
// Mirrors the *inputs* the reference hands to its compilers:
//   RowExpression IR  main/sql/relational/{CallExpression,SpecialForm,ConstantExpression,InputReferenceExpression}.java
//   aggregation spec  main/operator/HashAggregationOperator.java:57-135 (factory arguments)
// and rejects everything outside the supported subset with NOT_SUPPORTED at plan time.
#pragma once
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace pgpu {

typedef __int128 i128;
typedef unsigned __int128 u128;

enum class TypeKind : int { BOOLEAN = 0, BIGINT = 1, INTEGER = 2, DATE = 3, DOUBLE = 4, DECIMAL = 5, VARCHAR = 6, VARBINARY = 7 };

struct Type {
    TypeKind kind = TypeKind::BIGINT;
    int precision = 0;  // DECIMAL precision, VARCHAR bound (0 = unbounded)
    int scale = 0;
    bool operator==(const Type& o) const { return kind == o.kind && precision == o.precision && scale == o.scale; }
    bool operator!=(const Type& o) const { return !(*this == o); }
    bool isDecimal() const { return kind == TypeKind::DECIMAL; }
    bool isShortDecimal() const { return kind == TypeKind::DECIMAL && precision <= 18; }  // spi/type/Decimals.java MAX_SHORT_PRECISION
    bool isLongDecimal() const { return kind == TypeKind::DECIMAL && precision > 18; }
    bool isVarWidth() const { return kind == TypeKind::VARCHAR || kind == TypeKind::VARBINARY; }
    bool isInt32() const { return kind == TypeKind::INTEGER || kind == TypeKind::DATE; }
    bool isInt64Like() const { return kind == TypeKind::BIGINT || isShortDecimal(); }
    // bytes per position of the flat block; 0 for var-width
    int width() const;
    std::string str() const;
    static Type boolean() { return Type{TypeKind::BOOLEAN, 0, 0}; }
    static Type bigint() { return Type{TypeKind::BIGINT, 0, 0}; }
    static Type integer() { return Type{TypeKind::INTEGER, 0, 0}; }
    static Type date() { return Type{TypeKind::DATE, 0, 0}; }
    static Type dbl() { return Type{TypeKind::DOUBLE, 0, 0}; }
    static Type decimal(int p, int s) { return Type{TypeKind::DECIMAL, p, s}; }
    static Type varchar(int n) { return Type{TypeKind::VARCHAR, n, 0}; }
    static Type varbinary() { return Type{TypeKind::VARBINARY, 0, 0}; }
};

enum class ExprKind { INPUT, CONST, CALL, AND, OR, BETWEEN, IN, IS_NULL, IF, COALESCE };

// Operators of CALL nodes (io.trino.spi.function.OperatorType + $not + CAST).
enum class Op { ADD, SUBTRACT, MULTIPLY, DIVIDE, MODULUS, NEGATE, EQ, NE, LT, LE, GT, GE, NOT, CAST };

struct Expr {
    ExprKind kind = ExprKind::CONST;
    Type type;
    int channel = -1;        // INPUT
    bool isNull = false;     // CONST null
    int64_t i64 = 0;         // CONST: boolean/integer/date/bigint/short decimal value, or DOUBLE bits
    i128 dec = 0;            // CONST: long decimal (two's complement here; blocks are sign-magnitude)
    Op op = Op::ADD;         // CALL
    std::vector<std::unique_ptr<Expr>> args;
};

enum class AggFn { SUM, COUNT, AVG, MIN, MAX };
enum class Step { SINGLE, PARTIAL, FINAL, INTERMEDIATE };  // main/sql/planner/plan/AggregationNode.java:348-380 Step
// Step.isInputRaw() / isOutputPartial() of the reference, negated / as is
inline bool inputIsIntermediate(Step s) { return s == Step::FINAL || s == Step::INTERMEDIATE; }
inline bool outputIsIntermediate(Step s) { return s == Step::PARTIAL || s == Step::INTERMEDIATE; }

struct AggSpec {
    AggFn fn = AggFn::COUNT;
    Type outType;          // final output type
    Type inType;           // type of the raw argument (also needed in FINAL step to pick the state layout)
    bool hasArg = false;   // false: count(*)
    int arg = -1;          // raw step: channel of the argument; FINAL: first intermediate channel
    int mask = -1;         // optional boolean mask channel (main/operator/aggregation/AccumulatorCompiler.java:535-553)
};

struct PlanError : std::runtime_error {
    int status;  // pgpu_status
    PlanError(int st, const std::string& m) : std::runtime_error(m), status(st) {}
};

struct Plan {
    std::vector<Type> input;
    std::unique_ptr<Expr> filter;                     // may be null
    bool hasProject = false;
    std::vector<std::unique_ptr<Expr>> projections;   // identity INPUT refs when !hasProject
    bool hasAgg = false;
    Step step = Step::SINGLE;
    std::vector<int> groupChannels;                   // channels of the projected page
    int64_t expectedGroups = 10000;
    std::vector<AggSpec> aggs;
    // options
    int64_t maxPartialGroups = 0;   // PARTIAL: flush when the table holds this many groups (0 = never)
    int64_t maxOutputRows = 0;      // rows per output page (0 = one page)
    // filter + project only: MergePages (operator/project/MergePages.java:60-84). Output pages with fewer than mergeMinRows rows
    // and fewer than mergeMinBytes bytes are buffered; both 0 = every page goes out as it is (no merging)
    int64_t mergeMinRows = 0;       // session property filter_and_project_min_output_page_row_count (256 in the reference)
    int64_t mergeMinBytes = 0;      // filter_and_project_min_output_page_size (500 kB in the reference)
    int64_t mergeMaxBytes = 1 << 20;  // PageBuilder.DEFAULT_MAX_PAGE_SIZE_IN_BYTES: the buffer is flushed when it holds this much
    std::string text;               // canonical source text

    // schema of the page after projection (input to aggregation / output of filter-project)
    std::vector<Type> projectedTypes() const;
    // schema of what get_output returns
    std::vector<Type> outputTypes() const;
    // FINAL-step intermediate layout / PARTIAL-step output layout of one aggregate
    static std::vector<Type> intermediateTypes(const AggSpec& a);
};

// Parses and validates; throws PlanError (PGPU_INVALID_ARGUMENT / PGPU_NOT_SUPPORTED).
std::unique_ptr<Plan> parsePlan(const std::string& text);

// helpers shared with codegen / oracle
const char* opName(Op op);
const char* aggName(AggFn f);
i128 pow10_i128(int n);
std::string i128ToString(i128 v);
// Result type rules of main/type/DecimalOperators.java (add :84-85, multiply :261-262, ...); throws NOT_SUPPORTED when invalid.
Type decimalResultType(Op op, const Type& a, const Type& b);

}  // namespace pgpu
