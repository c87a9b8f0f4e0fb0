// plan.cpp — parser + validator for the plan language (see plan.h, DESIGN.md §"Plan language").
#include "plan.h"

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "../../include/prestogpu.h"

namespace pgpu {

int Type::width() const {
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

std::string Type::str() const {
    switch (kind) {
        case TypeKind::BOOLEAN: return "boolean";
        case TypeKind::BIGINT: return "bigint";
        case TypeKind::INTEGER: return "integer";
        case TypeKind::DATE: return "date";
        case TypeKind::DOUBLE: return "double";
        case TypeKind::DECIMAL: return "(decimal " + std::to_string(precision) + " " + std::to_string(scale) + ")";
        case TypeKind::VARCHAR: return precision > 0 ? "(varchar " + std::to_string(precision) + ")" : "varchar";
        case TypeKind::VARBINARY: return "varbinary";
    }
    return "?";
}

const char* opName(Op op) {
    switch (op) {
        case Op::ADD: return "add";
        case Op::SUBTRACT: return "subtract";
        case Op::MULTIPLY: return "multiply";
        case Op::DIVIDE: return "divide";
        case Op::MODULUS: return "modulus";
        case Op::NEGATE: return "negate";
        case Op::EQ: return "eq";
        case Op::NE: return "ne";
        case Op::LT: return "lt";
        case Op::LE: return "le";
        case Op::GT: return "gt";
        case Op::GE: return "ge";
        case Op::NOT: return "not";
        case Op::CAST: return "cast";
    }
    return "?";
}

const char* aggName(AggFn f) {
    switch (f) {
        case AggFn::SUM: return "sum";
        case AggFn::COUNT: return "count";
        case AggFn::AVG: return "avg";
        case AggFn::MIN: return "min";
        case AggFn::MAX: return "max";
    }
    return "?";
}

i128 pow10_i128(int n) {
    i128 v = 1;
    for (int i = 0; i < n; i++) v *= 10;
    return v;
}

std::string i128ToString(i128 v) {
    if (v == 0) return "0";
    bool neg = v < 0;
    u128 m = neg ? (u128)0 - (u128)v : (u128)v;
    std::string s;
    while (m) {
        s.push_back((char)('0' + (int)(m % 10)));
        m /= 10;
    }
    if (neg) s.push_back('-');
    std::reverse(s.begin(), s.end());
    return s;
}

namespace {

[[noreturn]] void bad(const std::string& m) { throw PlanError(PGPU_INVALID_ARGUMENT, "invalid plan: " + m); }
[[noreturn]] void unsupported(const std::string& m) { throw PlanError(PGPU_NOT_SUPPORTED, "not supported on the GPU path: " + m); }

// ---------------------------------------------------------------- S-expressions
struct Sx {
    bool atom = false;
    std::string text;
    std::vector<Sx> kids;
    const std::string& head() const {
        static const std::string empty;
        if (atom || kids.empty() || !kids[0].atom) return empty;
        return kids[0].text;
    }
};

struct Lexer {
    const std::string& s;
    size_t i = 0;
    explicit Lexer(const std::string& src) : s(src) {}
    void skip() {
        while (i < s.size()) {
            if (isspace((unsigned char)s[i])) i++;
            else if (s[i] == ';') { while (i < s.size() && s[i] != '\n') i++; }
            else break;
        }
    }
    Sx parse() {
        skip();
        if (i >= s.size()) bad("unexpected end of text");
        Sx x;
        if (s[i] == '(') {
            i++;
            for (;;) {
                skip();
                if (i >= s.size()) bad("missing ')'");
                if (s[i] == ')') { i++; break; }
                x.kids.push_back(parse());
            }
        } else if (s[i] == ')') {
            bad("unexpected ')'");
        } else {
            x.atom = true;
            size_t j = i;
            while (j < s.size() && !isspace((unsigned char)s[j]) && s[j] != '(' && s[j] != ')') j++;
            x.text = s.substr(i, j - i);
            i = j;
        }
        return x;
    }
};

int64_t parseInt(const Sx& x, const char* what) {
    if (!x.atom) bad(std::string("expected integer for ") + what);
    errno = 0;
    char* end = nullptr;
    long long v = strtoll(x.text.c_str(), &end, 0);
    if (errno || end == x.text.c_str() || *end) bad(std::string("bad integer '") + x.text + "' for " + what);
    return v;
}

i128 parseI128(const std::string& t) {
    size_t k = 0;
    bool neg = false;
    if (k < t.size() && (t[k] == '-' || t[k] == '+')) { neg = t[k] == '-'; k++; }
    if (k >= t.size()) bad("bad decimal literal '" + t + "'");
    u128 m = 0;
    for (; k < t.size(); k++) {
        if (!isdigit((unsigned char)t[k])) bad("bad decimal literal '" + t + "'");
        u128 nm = m * 10 + (unsigned)(t[k] - '0');
        if (nm / 10 != m) bad("decimal literal out of range '" + t + "'");
        m = nm;
    }
    if (m >> 127) bad("decimal literal out of range '" + t + "'");
    return neg ? -(i128)m : (i128)m;
}

Type parseType(const Sx& x) {
    if (x.atom) {
        const std::string& t = x.text;
        if (t == "boolean") return Type::boolean();
        if (t == "bigint") return Type::bigint();
        if (t == "integer") return Type::integer();
        if (t == "date") return Type::date();
        if (t == "double") return Type::dbl();
        if (t == "varchar") return Type::varchar(0);
        if (t == "varbinary") return Type::varbinary();
        if (t == "real" || t == "smallint" || t == "tinyint" || t == "timestamp" || t == "char" || t == "array" || t == "map" || t == "row")
            unsupported("type " + t);
        bad("unknown type '" + t + "'");
    }
    const std::string& h = x.head();
    if (h == "decimal") {
        if (x.kids.size() != 3) bad("(decimal p s)");
        int p = (int)parseInt(x.kids[1], "precision"), s = (int)parseInt(x.kids[2], "scale");
        if (p < 1 || p > 38 || s < 0 || s > p) bad("decimal precision/scale out of range");  // spi/type/DecimalType.java:42-57
        return Type::decimal(p, s);
    }
    if (h == "varchar") {
        if (x.kids.size() != 2) bad("(varchar n)");
        int n = (int)parseInt(x.kids[1], "varchar length");
        if (n < 0) bad("varchar length");
        return Type::varchar(n);
    }
    if (h == "array" || h == "map" || h == "row" || h == "char" || h == "timestamp") unsupported("type " + h);
    bad("unknown type form '" + h + "'");
}

bool isIntegral(const Type& t) { return t.kind == TypeKind::BIGINT || t.kind == TypeKind::INTEGER; }

Op parseOp(const std::string& t) {
    static const std::pair<const char*, Op> tbl[] = {
        {"add", Op::ADD}, {"subtract", Op::SUBTRACT}, {"multiply", Op::MULTIPLY}, {"divide", Op::DIVIDE},
        {"modulus", Op::MODULUS}, {"negate", Op::NEGATE}, {"eq", Op::EQ}, {"ne", Op::NE}, {"lt", Op::LT},
        {"le", Op::LE}, {"gt", Op::GT}, {"ge", Op::GE}, {"not", Op::NOT}, {"cast", Op::CAST}};
    for (auto& e : tbl)
        if (t == e.first) return e.second;
    unsupported("scalar function '" + t + "' (only operators, NOT and CAST compile to the GPU)");
}

struct ExprParser {
    const std::vector<Type>& input;
    explicit ExprParser(const std::vector<Type>& in) : input(in) {}

    std::unique_ptr<Expr> constant(const Sx& x) {
        if (x.kids.size() != 3) bad("(const <type> <literal>)");
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::CONST;
        e->type = parseType(x.kids[1]);
        if (!x.kids[2].atom) bad("constant literal");
        const std::string& lit = x.kids[2].text;
        if (lit == "null") { e->isNull = true; return e; }
        switch (e->type.kind) {
            case TypeKind::BOOLEAN:
                if (lit == "true") e->i64 = 1;
                else if (lit == "false") e->i64 = 0;
                else bad("boolean literal '" + lit + "'");
                break;
            case TypeKind::BIGINT: e->i64 = parseInt(x.kids[2], "bigint literal"); break;
            case TypeKind::INTEGER:
            case TypeKind::DATE: {
                int64_t v = parseInt(x.kids[2], "integer literal");
                if (v < INT32_MIN || v > INT32_MAX) bad("integer literal out of range");
                e->i64 = v;
                break;
            }
            case TypeKind::DOUBLE: {
                if (lit.size() > 2 && lit[0] == '0' && (lit[1] == 'x' || lit[1] == 'X') && lit.find('.') == std::string::npos && lit.find('p') == std::string::npos) {
                    // raw IEEE bits, exact
                    errno = 0;
                    unsigned long long b = strtoull(lit.c_str(), nullptr, 16);
                    if (errno) bad("double bits literal");
                    e->i64 = (int64_t)b;
                } else {
                    char* end = nullptr;
                    double d = strtod(lit.c_str(), &end);
                    if (end == lit.c_str() || *end) bad("double literal '" + lit + "'");
                    memcpy(&e->i64, &d, 8);
                }
                break;
            }
            case TypeKind::DECIMAL: {
                i128 v = parseI128(lit);
                i128 lim = pow10_i128(e->type.precision);
                if (v >= lim || v <= -lim) bad("decimal literal exceeds precision");
                e->dec = v;
                e->i64 = (int64_t)v;
                break;
            }
            default: unsupported("constant of type " + e->type.str());
        }
        return e;
    }

    std::unique_ptr<Expr> parse(const Sx& x) {
        if (x.atom) bad("expected expression, got '" + x.text + "'");
        const std::string& h = x.head();
        auto e = std::make_unique<Expr>();
        if (h == "in") {
            if (x.kids.size() != 3) bad("(in <channel> <type>)");
            e->kind = ExprKind::INPUT;
            e->channel = (int)parseInt(x.kids[1], "channel");
            e->type = parseType(x.kids[2]);
            if (e->channel < 0 || e->channel >= (int)input.size()) bad("input channel " + std::to_string(e->channel) + " out of range");
            if (input[e->channel] != e->type) bad("input channel " + std::to_string(e->channel) + " is " + input[e->channel].str() + ", expression says " + e->type.str());
            return e;
        }
        if (h == "const") return constant(x);
        if (h == "call") {
            if (x.kids.size() < 4) bad("(call <op> <type> <args>...)");
            if (!x.kids[1].atom) bad("call operator name");
            e->kind = ExprKind::CALL;
            e->op = parseOp(x.kids[1].text);
            e->type = parseType(x.kids[2]);
            for (size_t k = 3; k < x.kids.size(); k++) e->args.push_back(parse(x.kids[k]));
            checkCall(*e);
            return e;
        }
        if (h == "and" || h == "or") {
            if (x.kids.size() < 3) bad("(" + h + " a b ...)");
            e->kind = h == "and" ? ExprKind::AND : ExprKind::OR;
            e->type = Type::boolean();
            for (size_t k = 1; k < x.kids.size(); k++) {
                e->args.push_back(parse(x.kids[k]));
                if (e->args.back()->type.kind != TypeKind::BOOLEAN) bad(h + " operand must be boolean");
            }
            return e;
        }
        if (h == "between") {  // main/sql/gen/BetweenCodeGenerator.java: value >= min AND value <= max
            if (x.kids.size() != 4) bad("(between v lo hi)");
            e->kind = ExprKind::BETWEEN;
            e->type = Type::boolean();
            for (size_t k = 1; k < 4; k++) e->args.push_back(parse(x.kids[k]));
            if (e->args[0]->type != e->args[1]->type || e->args[0]->type != e->args[2]->type) bad("between operand types differ");
            checkComparable(e->args[0]->type, true);
            return e;
        }
        if (h == "in-list") {  // main/sql/gen/InCodeGenerator.java
            if (x.kids.size() < 3) bad("(in-list v c...)");
            e->kind = ExprKind::IN;
            e->type = Type::boolean();
            for (size_t k = 1; k < x.kids.size(); k++) e->args.push_back(parse(x.kids[k]));
            for (size_t k = 1; k < e->args.size(); k++) {
                if (e->args[k]->type != e->args[0]->type) bad("in-list operand types differ");
                if (e->args[k]->kind != ExprKind::CONST) unsupported("IN list with non-constant members");
            }
            checkComparable(e->args[0]->type, false);
            return e;
        }
        if (h == "is-null") {
            if (x.kids.size() != 2) bad("(is-null e)");
            e->kind = ExprKind::IS_NULL;
            e->type = Type::boolean();
            e->args.push_back(parse(x.kids[1]));
            return e;
        }
        if (h == "if") {  // main/sql/gen/IfCodeGenerator.java
            if (x.kids.size() != 5) bad("(if <type> c a b)");
            e->kind = ExprKind::IF;
            e->type = parseType(x.kids[1]);
            for (size_t k = 2; k < 5; k++) e->args.push_back(parse(x.kids[k]));
            if (e->args[0]->type.kind != TypeKind::BOOLEAN) bad("if condition must be boolean");
            if (e->args[1]->type != e->type || e->args[2]->type != e->type) bad("if branch types differ from result type");
            checkValueType(e->type);
            return e;
        }
        if (h == "coalesce") {  // main/sql/gen/CoalesceCodeGenerator.java
            if (x.kids.size() < 4) bad("(coalesce <type> a b ...)");
            e->kind = ExprKind::COALESCE;
            e->type = parseType(x.kids[1]);
            for (size_t k = 2; k < x.kids.size(); k++) {
                e->args.push_back(parse(x.kids[k]));
                if (e->args.back()->type != e->type) bad("coalesce operand type differs from result type");
            }
            checkValueType(e->type);
            return e;
        }
        if (h == "null-if") {
            // main/sql/gen/NullIfCodeGenerator.java:70-113: first NULL -> NULL without evaluating second; equal(first, second)
            // true -> NULL; false or NULL -> first. Lowered to IF(IS_NULL(a), NULL, IF(EQ(a, b), NULL, a)); IF evaluates only the
            // branch a row takes and treats a NULL condition as false, which is exactly that control flow. The casts to a common
            // type the reference carries as function dependencies are explicit (call cast ...) nodes in this plan language.
            if (x.kids.size() != 3) bad("(null-if a b)");
            auto a = parse(x.kids[1]);
            auto b = parse(x.kids[2]);
            if (a->type != b->type) bad("null-if operand types differ (wrap one side in (call cast ...)): " + a->type.str() + " vs " + b->type.str());
            const Type t = a->type;
            checkValueType(t);
            checkComparable(t, false);
            auto inner = makeIf(t, makeEq(std::move(a), std::move(b)), makeNull(t), parse(x.kids[1]));
            return makeIf(t, makeIsNull(parse(x.kids[1])), makeNull(t), std::move(inner));
        }
        if (h == "switch") {
            // main/sql/gen/SwitchCodeGenerator.java:113-159: value evaluated once, NULL value -> else branch; otherwise the first
            // WHEN whose equal(operand, value) is true (a NULL comparison counts as false) yields its result; no ELSE -> NULL.
            // Lowered to IF(IS_NULL(v), else, IF(EQ(o1, v), r1, IF(EQ(o2, v), r2, ... else))). (Expressions are deterministic, so
            // re-evaluating the value per comparison is unobservable.)
            if (x.kids.size() < 4) bad("(switch <type> v (when o r)... [else])");
            const Type t = parseType(x.kids[1]);
            checkValueType(t);
            size_t last = x.kids.size();  // one past the last WHEN clause
            const Sx* elseSx = nullptr;
            if (x.kids.back().atom || x.kids.back().head() != "when") {
                elseSx = &x.kids.back();
                last--;
            }
            if (last < 4) bad("switch needs at least one (when operand result) clause");
            auto elseExpr = [&]() -> std::unique_ptr<Expr> {
                if (!elseSx) return makeNull(t);
                auto r = parse(*elseSx);
                if (r->type != t) bad("switch else type differs from result type");
                return r;
            };
            const Type vt = parse(x.kids[2])->type;
            checkComparable(vt, false);
            std::unique_ptr<Expr> chain = elseExpr();
            for (size_t k = last; k-- > 3;) {
                const Sx& w = x.kids[k];
                if (w.atom || w.head() != "when" || w.kids.size() != 3) bad("(when operand result)");
                auto o = parse(w.kids[1]);
                auto r = parse(w.kids[2]);
                if (o->type != vt) bad("when operand type differs from the switch value (wrap it in (call cast ...)): " + o->type.str() + " vs " + vt.str());
                if (r->type != t) bad("when result type differs from result type");
                chain = makeIf(t, makeEq(std::move(o), parse(x.kids[2])), std::move(r), std::move(chain));
            }
            return makeIf(t, makeIsNull(parse(x.kids[2])), elseExpr(), std::move(chain));
        }
        if (h == "when") bad("(when ...) outside (switch ...)");
        if (h == "dereference" || h == "row-constructor" || h == "bind" || h == "lambda")
            unsupported("special form " + h);
        bad("unknown expression form '" + h + "'");
    }

    // nodes of the lowered forms (null-if, switch)
    static std::unique_ptr<Expr> makeNull(const Type& t) {
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::CONST;
        e->type = t;
        e->isNull = true;
        return e;
    }
    static std::unique_ptr<Expr> makeIsNull(std::unique_ptr<Expr> a) {
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::IS_NULL;
        e->type = Type::boolean();
        e->args.push_back(std::move(a));
        return e;
    }
    static std::unique_ptr<Expr> makeEq(std::unique_ptr<Expr> a, std::unique_ptr<Expr> b) {
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::CALL;
        e->op = Op::EQ;
        e->type = Type::boolean();
        e->args.push_back(std::move(a));
        e->args.push_back(std::move(b));
        return e;
    }
    static std::unique_ptr<Expr> makeIf(const Type& t, std::unique_ptr<Expr> c, std::unique_ptr<Expr> a, std::unique_ptr<Expr> b) {
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::IF;
        e->type = t;
        e->args.push_back(std::move(c));
        e->args.push_back(std::move(a));
        e->args.push_back(std::move(b));
        return e;
    }

    static void checkValueType(const Type& t) {
        if (t.isVarWidth()) unsupported("computing " + t.str() + " values in expressions (var-width columns can only be passed through or grouped on)");
    }

    static void checkComparable(const Type& t, bool ordered) {
        switch (t.kind) {
            case TypeKind::BOOLEAN:
                if (ordered) unsupported("ordering comparison on boolean");
                return;
            case TypeKind::BIGINT:
            case TypeKind::INTEGER:
            case TypeKind::DATE:
            case TypeKind::DOUBLE:
            case TypeKind::DECIMAL: return;
            default: unsupported("comparison on " + t.str());
        }
    }

    void checkCall(const Expr& e) {
        auto argc = [&](size_t n) {
            if (e.args.size() != n) bad(std::string(opName(e.op)) + " takes " + std::to_string(n) + " argument(s)");
        };
        switch (e.op) {
            case Op::ADD:
            case Op::SUBTRACT:
            case Op::MULTIPLY:
            case Op::DIVIDE:
            case Op::MODULUS: {
                argc(2);
                const Type &a = e.args[0]->type, &b = e.args[1]->type;
                if (a.kind == TypeKind::DECIMAL && b.kind == TypeKind::DECIMAL) {
                    Type r = decimalResultType(e.op, a, b);
                    if (r != e.type) bad(std::string("decimal ") + opName(e.op) + " of " + a.str() + ", " + b.str() + " yields " + r.str() + ", plan says " + e.type.str());
                    return;
                }
                if (a != b || a != e.type) bad(std::string(opName(e.op)) + " operand/result types must match (" + a.str() + ", " + b.str() + " -> " + e.type.str() + ")");
                if (a.kind == TypeKind::BIGINT || a.kind == TypeKind::INTEGER || a.kind == TypeKind::DOUBLE) return;
                unsupported(std::string(opName(e.op)) + " on " + a.str());
            }
            case Op::NEGATE: {
                argc(1);
                const Type& a = e.args[0]->type;
                if (a != e.type) bad("negate result type");
                if (a.kind == TypeKind::BIGINT || a.kind == TypeKind::INTEGER || a.kind == TypeKind::DOUBLE || a.kind == TypeKind::DECIMAL) return;
                unsupported("negate on " + a.str());
            }
            case Op::EQ:
            case Op::NE:
            case Op::LT:
            case Op::LE:
            case Op::GT:
            case Op::GE: {
                argc(2);
                if (e.type.kind != TypeKind::BOOLEAN) bad("comparison must return boolean");
                if (e.args[0]->type != e.args[1]->type) bad("comparison operand types differ: " + e.args[0]->type.str() + " vs " + e.args[1]->type.str());
                checkComparable(e.args[0]->type, !(e.op == Op::EQ || e.op == Op::NE));
                return;
            }
            case Op::NOT:
                argc(1);
                if (e.type.kind != TypeKind::BOOLEAN || e.args[0]->type.kind != TypeKind::BOOLEAN) bad("not is boolean -> boolean");
                return;
            case Op::CAST: {
                argc(1);
                const Type &f = e.args[0]->type, &t = e.type;
                if (f == t) return;
                auto num = [](const Type& x) { return x.kind == TypeKind::BIGINT || x.kind == TypeKind::INTEGER || x.kind == TypeKind::DOUBLE || x.kind == TypeKind::DECIMAL; };
                if (num(f) && num(t)) return;
                if (f.kind == TypeKind::BOOLEAN && (isIntegral(t) || t.kind == TypeKind::DOUBLE)) return;
                if ((isIntegral(f) || f.kind == TypeKind::DOUBLE) && t.kind == TypeKind::BOOLEAN) return;
                unsupported("cast " + f.str() + " -> " + t.str());
            }
        }
    }
};

AggFn parseAggFn(const std::string& t) {
    if (t == "sum") return AggFn::SUM;
    if (t == "count") return AggFn::COUNT;
    if (t == "avg") return AggFn::AVG;
    if (t == "min") return AggFn::MIN;
    if (t == "max") return AggFn::MAX;
    unsupported("aggregate function '" + t + "' (GPU path has sum/count/avg/min/max)");
}

// Return type rules: DoubleSumAggregation/LongSumAggregation (same type), DecimalSumAggregation.java:76-77
// (decimal(38,s)), AverageAggregations (double), DecimalAverageAggregation (same decimal type), min/max (same type).
void checkAgg(const AggSpec& a) {
    const Type& in = a.inType;
    auto no = [&]() { unsupported(std::string(aggName(a.fn)) + "(" + in.str() + ")"); };
    switch (a.fn) {
        case AggFn::COUNT:
            if (a.outType.kind != TypeKind::BIGINT) bad("count returns bigint");
            return;
        case AggFn::SUM:
            if (in.kind == TypeKind::BIGINT || in.kind == TypeKind::DOUBLE) {
                if (a.outType != in) bad("sum(" + in.str() + ") returns " + in.str());
                return;
            }
            if (in.kind == TypeKind::DECIMAL) {
                if (a.outType != Type::decimal(38, in.scale)) bad("sum(" + in.str() + ") returns decimal(38," + std::to_string(in.scale) + ")");
                return;
            }
            no();
            return;
        case AggFn::AVG:
            if (in.kind == TypeKind::BIGINT || in.kind == TypeKind::DOUBLE) {
                if (a.outType.kind != TypeKind::DOUBLE) bad("avg(" + in.str() + ") returns double");
                return;
            }
            if (in.kind == TypeKind::DECIMAL) {
                if (a.outType != in) bad("avg(" + in.str() + ") returns " + in.str());
                return;
            }
            no();
            return;
        case AggFn::MIN:
        case AggFn::MAX:
            if (a.outType != in) bad("min/max return their argument type");
            if (in.kind == TypeKind::BIGINT || in.kind == TypeKind::INTEGER || in.kind == TypeKind::DATE || in.kind == TypeKind::DOUBLE ||
                in.kind == TypeKind::DECIMAL || in.kind == TypeKind::BOOLEAN)
                return;
            // VARCHAR(n <= 15): the value fits the 16-byte cell the GPU path keeps varchar values in (one 128-bit min / max cell)
            if (in.kind == TypeKind::VARCHAR && in.precision >= 1 && in.precision <= 15) return;
            no();
            return;
    }
}

}  // namespace

Type decimalResultType(Op op, const Type& a, const Type& b) {
    int ap = a.precision, as = a.scale, bp = b.precision, bs = b.scale;
    int p = 0, s = 0;
    switch (op) {
        case Op::ADD:
        case Op::SUBTRACT:  // main/type/DecimalOperators.java:84-85,186-187
            s = std::max(as, bs);
            p = std::min(38, std::max(ap - as, bp - bs) + s + 1);
            break;
        case Op::MULTIPLY:  // :261-262
            s = as + bs;
            p = std::min(38, ap + bp);
            if (s > 38) unsupported("decimal multiply with result scale > 38");
            break;
        case Op::DIVIDE:  // :339-340
            s = std::max(as, bs);
            p = std::min(38, ap + bs + std::max(0, bs - as));
            break;
        case Op::MODULUS:  // :452-453
            s = std::max(as, bs);
            p = std::min(bp - bs, ap - as) + s;
            break;
        default: bad("not a decimal arithmetic operator");
    }
    if (p < 1) p = 1;
    return Type::decimal(p, s);
}

std::vector<Type> Plan::intermediateTypes(const AggSpec& a) {
    // Flattened field order of the reference's serialized states (state/StateCompiler.java:588-630: primitives
    // first, then by field name): LongDoubleState(first,second), LongLongState(first,second),
    // LongAndDoubleState(double,long), LongState, Nullable{Long,Double}State; decimal states are VARBINARY
    // (state/LongDecimalWithOverflowStateSerializer.java:43-50, LongDecimalWithOverflowAndLongStateSerializer.java:43-52).
    switch (a.fn) {
        case AggFn::COUNT: return {Type::bigint()};
        case AggFn::SUM:
            if (a.inType.kind == TypeKind::DOUBLE) return {Type::bigint(), Type::dbl()};
            if (a.inType.kind == TypeKind::BIGINT) return {Type::bigint(), Type::bigint()};
            return {Type::varbinary()};
        case AggFn::AVG:
            if (a.inType.kind == TypeKind::DECIMAL) return {Type::varbinary()};
            return {Type::dbl(), Type::bigint()};
        case AggFn::MIN:
        case AggFn::MAX: return {a.inType};
    }
    return {};
}

std::vector<Type> Plan::projectedTypes() const {
    std::vector<Type> t;
    for (auto& p : projections) t.push_back(p->type);
    return t;
}

std::vector<Type> Plan::outputTypes() const {
    std::vector<Type> proj = projectedTypes();
    if (!hasAgg) return proj;
    std::vector<Type> out;
    for (int g : groupChannels) out.push_back(proj[g]);
    for (auto& a : aggs) {
        if (outputIsIntermediate(step)) {
            for (auto& t : intermediateTypes(a)) out.push_back(t);
        } else {
            out.push_back(a.outType);
        }
    }
    return out;
}

std::unique_ptr<Plan> parsePlan(const std::string& text) {
    Lexer lx(text);
    Sx root = lx.parse();
    lx.skip();
    if (lx.i != text.size()) bad("trailing text after plan");
    if (root.head() != "plan") bad("expected (plan ...)");
    auto plan = std::make_unique<Plan>();
    plan->text = text;
    const Sx *sInput = nullptr, *sFilter = nullptr, *sProject = nullptr, *sAgg = nullptr, *sOpt = nullptr;
    for (size_t k = 1; k < root.kids.size(); k++) {
        const std::string& h = root.kids[k].head();
        if (h == "input") sInput = &root.kids[k];
        else if (h == "filter") sFilter = &root.kids[k];
        else if (h == "project") sProject = &root.kids[k];
        else if (h == "aggregate") sAgg = &root.kids[k];
        else if (h == "options") sOpt = &root.kids[k];
        else bad("unknown plan section '" + h + "'");
    }
    if (!sInput) bad("missing (input ...)");
    for (size_t k = 1; k < sInput->kids.size(); k++) plan->input.push_back(parseType(sInput->kids[k]));
    if (plan->input.empty()) bad("plan has no input channels");
    if (plan->input.size() > 64) unsupported("more than 64 input channels");
    ExprParser ep(plan->input);
    if (sFilter) {
        if (sFilter->kids.size() != 2) bad("(filter <expr>)");
        plan->filter = ep.parse(sFilter->kids[1]);
        if (plan->filter->type.kind != TypeKind::BOOLEAN) bad("filter must be boolean");
    }
    if (sProject) {
        plan->hasProject = true;
        for (size_t k = 1; k < sProject->kids.size(); k++) plan->projections.push_back(ep.parse(sProject->kids[k]));
        if (plan->projections.empty()) bad("(project) needs at least one expression");
    } else {
        for (size_t c = 0; c < plan->input.size(); c++) {
            auto e = std::make_unique<Expr>();
            e->kind = ExprKind::INPUT;
            e->channel = (int)c;
            e->type = plan->input[c];
            plan->projections.push_back(std::move(e));
        }
    }
    for (auto& p : plan->projections) {
        if (p->type.isVarWidth() && p->kind != ExprKind::INPUT) unsupported("var-width projection other than a column reference");
    }
    std::vector<Type> proj = plan->projectedTypes();
    if (sAgg) {
        plan->hasAgg = true;
        for (size_t k = 1; k < sAgg->kids.size(); k++) {
            const Sx& s = sAgg->kids[k];
            const std::string& h = s.head();
            if (h == "step") {
                if (s.kids.size() != 2 || !s.kids[1].atom) bad("(step single|partial|final)");
                const std::string& t = s.kids[1].text;
                if (t == "single") plan->step = Step::SINGLE;
                else if (t == "partial") plan->step = Step::PARTIAL;
                else if (t == "final") plan->step = Step::FINAL;
                else if (t == "intermediate") plan->step = Step::INTERMEDIATE;
                else bad("unknown step '" + t + "'");
            } else if (h == "group") {
                for (size_t j = 1; j < s.kids.size(); j++) {
                    int c = (int)parseInt(s.kids[j], "group channel");
                    if (c < 0 || c >= (int)proj.size()) bad("group channel out of range");
                    plan->groupChannels.push_back(c);
                }
            } else if (h == "expected-groups") {
                if (s.kids.size() != 2) bad("(expected-groups n)");
                plan->expectedGroups = parseInt(s.kids[1], "expected-groups");
                if (plan->expectedGroups < 1) plan->expectedGroups = 1;
            } else if (h == "agg") {
                // (agg <fn> <out-type> [(arg <ch> <in-type>)] [(mask <ch>)])
                if (s.kids.size() < 3 || !s.kids[1].atom) bad("(agg fn out-type ...)");
                AggSpec a;
                a.fn = parseAggFn(s.kids[1].text);
                a.outType = parseType(s.kids[2]);
                for (size_t j = 3; j < s.kids.size(); j++) {
                    const std::string& ah = s.kids[j].head();
                    if (ah == "arg") {
                        if (s.kids[j].kids.size() != 3) bad("(arg <channel> <type>)");
                        a.hasArg = true;
                        a.arg = (int)parseInt(s.kids[j].kids[1], "agg channel");
                        a.inType = parseType(s.kids[j].kids[2]);
                    } else if (ah == "mask") {
                        if (s.kids[j].kids.size() != 2) bad("(mask <channel>)");
                        a.mask = (int)parseInt(s.kids[j].kids[1], "mask channel");
                    } else if (ah == "distinct" || ah == "order-by") {
                        unsupported("aggregate with " + ah);
                    } else bad("unknown agg clause '" + ah + "'");
                }
                if (!a.hasArg) {
                    if (a.fn != AggFn::COUNT) bad(std::string(aggName(a.fn)) + " needs (arg ...)");
                    a.inType = Type::bigint();
                }
                checkAgg(a);
                plan->aggs.push_back(a);
            } else bad("unknown aggregate clause '" + h + "'");
        }
        if (plan->aggs.empty() && plan->groupChannels.empty()) bad("aggregate without groups or functions");
        if (plan->aggs.size() > 32) unsupported("more than 32 aggregates");
        // channel checks
        for (auto& a : plan->aggs) {
            if (inputIsIntermediate(plan->step)) {
                if (a.mask >= 0) bad("mask on FINAL step");
                std::vector<Type> it = Plan::intermediateTypes(a);
                int first = a.hasArg ? a.arg : -1;
                if (first < 0) bad("FINAL step aggregate needs (arg <first intermediate channel> <raw input type>)");
                for (size_t j = 0; j < it.size(); j++) {
                    int c = first + (int)j;
                    if (c >= (int)proj.size()) bad("intermediate channel out of range");
                    if (proj[c] != it[j]) bad("intermediate channel " + std::to_string(c) + " is " + proj[c].str() + ", state needs " + it[j].str());
                }
            } else {
                if (a.hasArg) {
                    if (a.arg < 0 || a.arg >= (int)proj.size()) bad("aggregate channel out of range");
                    if (a.fn != AggFn::COUNT && proj[a.arg] != a.inType) bad("aggregate channel " + std::to_string(a.arg) + " is " + proj[a.arg].str() + ", agg says " + a.inType.str());
                    if (a.fn == AggFn::COUNT) a.inType = proj[a.arg];
                }
                if (a.mask >= 0) {
                    if (a.mask >= (int)proj.size() || proj[a.mask].kind != TypeKind::BOOLEAN) bad("mask channel must be boolean");
                }
            }
        }
        for (int g : plan->groupChannels) {
            const Type& t = proj[g];
            if (t.kind == TypeKind::VARBINARY) unsupported("grouping on varbinary");
            if (t.kind == TypeKind::VARCHAR && (t.precision == 0 || t.precision > 15))
                unsupported("grouping on " + t.str() + " (GPU keys are fixed-width: varchar(n) with n <= 15)");
        }
        // var-width projections are only legal as group keys when aggregating
        for (size_t c = 0; c < proj.size(); c++) {
            if (!proj[c].isVarWidth()) continue;
            bool isKey = std::find(plan->groupChannels.begin(), plan->groupChannels.end(), (int)c) != plan->groupChannels.end();
            bool isState = false;
            if (inputIsIntermediate(plan->step))
                for (auto& a : plan->aggs) isState |= (a.arg == (int)c && proj[c].kind == TypeKind::VARBINARY);
            (void)isKey; (void)isState;  // unused var-width channels are simply ignored by the aggregation
        }
    } else {
        for (auto& t : proj)
            if (t.kind == TypeKind::VARBINARY) unsupported("varbinary output from filter/project");
    }
    if (sOpt) {
        for (size_t k = 1; k < sOpt->kids.size(); k++) {
            const Sx& s = sOpt->kids[k];
            const std::string& h = s.head();
            if (s.kids.size() != 2) bad("(options (key value)...)");
            if (h == "max-partial-groups") plan->maxPartialGroups = parseInt(s.kids[1], h.c_str());
            else if (h == "max-output-rows") plan->maxOutputRows = parseInt(s.kids[1], h.c_str());
            else if (h == "merge-min-rows") plan->mergeMinRows = parseInt(s.kids[1], h.c_str());
            else if (h == "merge-min-bytes") plan->mergeMinBytes = parseInt(s.kids[1], h.c_str());
            else if (h == "merge-max-bytes") plan->mergeMaxBytes = parseInt(s.kids[1], h.c_str());
            else bad("unknown option '" + h + "'");
        }
    }
    // MergePages.java:100-105, the same argument checks
    if (plan->mergeMinRows < 0 || plan->mergeMinBytes < 0) bad("merge-min-rows / merge-min-bytes must be greater or equal than zero");
    if (plan->mergeMaxBytes <= 0) bad("merge-max-bytes must be greater than zero");
    if (plan->mergeMaxBytes < plan->mergeMinBytes) bad("merge-max-bytes must be greater or equal than merge-min-bytes");
    if ((plan->mergeMinRows || plan->mergeMinBytes) && plan->hasAgg) bad("merge-min-rows / merge-min-bytes apply to filter + project plans only");
    return plan;
}

}  // namespace pgpu
