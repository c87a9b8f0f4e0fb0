// oracle/oracle_plan_expr.h — TEST INFRASTRUCTURE: second half of oracle_plan.h (expressions, aggregates, sections).
#pragma once

namespace oracle {
namespace plan_reader {

struct ExprReader {
    const std::vector<Type>& input;

    static Op opOf(const std::string& w) {
        static const char* names[] = {"add", "subtract", "multiply", "divide", "modulus", "negate", "eq", "ne", "lt", "le", "gt", "ge", "not", "cast"};
        for (int k = 0; k < 14; k++)
            if (w == names[k]) return (Op)k;  // same order as enum Op
        throw PlanError(kNotSupported, "oracle: scalar function '" + w + "' is outside the restated path");
    }

    std::unique_ptr<Expr> literal(Cursor& c) {
        auto e = std::make_unique<Expr>();
        e->kind = ExprKind::CONST;
        e->type = readType(c);
        const std::string lit = c.atom();
        if (lit == "null") {
            e->isNull = true;
            return e;
        }
        const TypeKind k = e->type.kind;
        if (k == TypeKind::BOOLEAN) {
            if (lit != "true" && lit != "false") malformed("boolean literal '" + lit + "'");
            e->i64 = lit == "true";
        } else if (k == TypeKind::DOUBLE) {
            const bool bits = lit.size() > 2 && lit[0] == '0' && (lit[1] == 'x' || lit[1] == 'X') && lit.find_first_of(".pP") == std::string::npos;
            if (bits) {  // the IEEE bit pattern, exact (relational.py writes doubles this way)
                e->i64 = (int64_t)strtoull(lit.c_str(), nullptr, 16);
            } else {
                char* end = nullptr;
                const double d = strtod(lit.c_str(), &end);
                if (end == lit.c_str() || *end) malformed("double literal '" + lit + "'");
                memcpy(&e->i64, &d, 8);
            }
        } else if (k == TypeKind::DECIMAL) {
            e->dec = readUnscaled(lit);
            const i128 bound = pow10_i128(e->type.precision);
            if (e->dec >= bound || e->dec <= -bound) malformed("decimal literal does not fit " + e->type.str());
            e->i64 = (int64_t)e->dec;
        } else if (k == TypeKind::BIGINT || k == TypeKind::INTEGER || k == TypeKind::DATE) {
            errno = 0;
            char* end = nullptr;
            const long long v = strtoll(lit.c_str(), &end, 0);
            if (errno || end == lit.c_str() || *end) malformed("integer literal '" + lit + "'");
            if (k != TypeKind::BIGINT && (v < INT_MIN || v > INT_MAX)) malformed("integer literal out of range");
            e->i64 = v;
        } else {
            throw PlanError(kNotSupported, "oracle: constant of type " + e->type.str());
        }
        return e;
    }

    // every operand up to the closing parenthesis
    void operands(Cursor& c, Expr& e) {
        while (!c.atClose()) e.args.push_back(read(c));
    }

    std::unique_ptr<Expr> read(Cursor& c) {
        const std::string form = c.enterNamed();
        std::unique_ptr<Expr> e;
        if (form == "const") {
            e = literal(c);
        } else {
            e = std::make_unique<Expr>();
            if (form == "in") {
                e->kind = ExprKind::INPUT;
                e->channel = (int)c.integer("input channel");
                e->type = readType(c);
                if (e->channel < 0 || e->channel >= (int)input.size()) malformed("input channel " + std::to_string(e->channel));
                if (input[(size_t)e->channel] != e->type) malformed("channel " + std::to_string(e->channel) + " is " + input[(size_t)e->channel].str() + ", not " + e->type.str());
            } else if (form == "call") {
                e->kind = ExprKind::CALL;
                e->op = opOf(c.atom());
                e->type = readType(c);
                operands(c, *e);
                const size_t want = e->op == Op::NEGATE || e->op == Op::NOT || e->op == Op::CAST ? 1 : 2;
                if (e->args.size() != want) malformed("operator with " + std::to_string(e->args.size()) + " operands");
            } else if (form == "and" || form == "or") {
                e->kind = form == "and" ? ExprKind::AND : ExprKind::OR;
                e->type = Type::boolean();
                operands(c, *e);
                if (e->args.size() < 2) malformed(form + " needs two operands");
            } else if (form == "between") {
                e->kind = ExprKind::BETWEEN;
                e->type = Type::boolean();
                operands(c, *e);
                if (e->args.size() != 3) malformed("(between value min max)");
            } else if (form == "in-list") {
                e->kind = ExprKind::IN;
                e->type = Type::boolean();
                operands(c, *e);
                if (e->args.size() < 2) malformed("(in-list value member...)");
                for (size_t k = 1; k < e->args.size(); k++)
                    if (e->args[k]->kind != ExprKind::CONST) throw PlanError(kNotSupported, "oracle: IN list with non-constant members");
            } else if (form == "is-null") {
                e->kind = ExprKind::IS_NULL;
                e->type = Type::boolean();
                operands(c, *e);
                if (e->args.size() != 1) malformed("(is-null e)");
            } else if (form == "if") {
                e->kind = ExprKind::IF;
                e->type = readType(c);
                operands(c, *e);
                if (e->args.size() != 3) malformed("(if type condition then else)");
            } else if (form == "coalesce") {
                e->kind = ExprKind::COALESCE;
                e->type = readType(c);
                operands(c, *e);
                if (e->args.size() < 2) malformed("(coalesce type a b ...)");
            } else if (form == "null-if") {  // Form.NULL_IF: the result has the type of the first operand (NullIfCodeGenerator.java:70-113)
                e->kind = ExprKind::NULL_IF;
                operands(c, *e);
                if (e->args.size() != 2) malformed("(null-if a b)");
                e->type = e->args[0]->type;
            } else if (form == "switch") {  // Form.SWITCH with Form.WHEN operands (SwitchCodeGenerator.java:113-159)
                e->kind = ExprKind::SWITCH;
                e->type = readType(c);
                e->args.push_back(read(c));
                while (!c.atClose()) {
                    if (c.peekName() == "when") {
                        if (e->hasElse) malformed("(when ...) after the else expression");
                        c.enterNamed();
                        e->args.push_back(read(c));
                        e->args.push_back(read(c));
                        c.leave();
                    } else {
                        if (e->hasElse) malformed("two else expressions");
                        e->hasElse = true;
                        e->args.push_back(read(c));
                    }
                }
                if (e->args.size() < 3) malformed("switch without a when clause");
            } else {
                throw PlanError(form == "when" ? kInvalidArgument : kNotSupported, "oracle: expression form '" + form + "'");
            }
        }
        c.leave();
        return e;
    }
};

inline AggSpec readAgg(Cursor& c) {  // after "(agg": fn out-type [(arg ch type)] [(mask ch)]
    AggSpec a;
    const std::string fn = c.atom();
    static const char* names[] = {"sum", "count", "avg", "min", "max"};  // enum AggFn order
    int which = -1;
    for (int k = 0; k < 5; k++)
        if (fn == names[k]) which = k;
    if (which < 0) throw PlanError(kNotSupported, "oracle: aggregate function '" + fn + "' is outside the restated path");
    a.fn = (AggFn)which;
    a.outType = readType(c);
    a.inType = Type::bigint();
    while (!c.atClose()) {
        const std::string clause = c.enterNamed();
        if (clause == "arg") {
            a.hasArg = true;
            a.arg = (int)c.integer("aggregate channel");
            a.inType = readType(c);
        } else if (clause == "mask") {
            a.mask = (int)c.integer("mask channel");
        } else {
            throw PlanError(kNotSupported, "oracle: aggregate clause '" + clause + "'");
        }
        c.leave();
    }
    if (!a.hasArg && a.fn != AggFn::COUNT) malformed(fn + " without an argument");
    return a;
}

}  // namespace plan_reader

inline std::unique_ptr<Plan> readPlan(const std::string& text) {
    using namespace plan_reader;
    Cursor c(text);
    if (c.enterNamed() != "plan") malformed("(plan ...) expected");
    auto plan = std::make_unique<Plan>();
    bool sawInput = false, sawProject = false;
    // sections may come in any order, but expressions need the input types: first pass finds (input ...)
    {
        Cursor scan(text);
        scan.enterNamed();
        while (!scan.atClose()) {
            if (scan.peekName() == "input") {
                scan.enterNamed();
                while (!scan.atClose()) plan->input.push_back(readType(scan));
                scan.leave();
                sawInput = true;
            } else {
                scan.skipList();
            }
        }
    }
    if (!sawInput || plan->input.empty()) malformed("no (input ...) section");
    ExprReader er{plan->input};
    while (!c.atClose()) {
        const std::string section = c.peekName();
        if (section == "input") {
            c.skipList();
            continue;
        }
        c.enterNamed();
        if (section == "filter") {
            plan->filter = er.read(c);
            if (plan->filter->type.kind != TypeKind::BOOLEAN) malformed("filter is " + plan->filter->type.str());
        } else if (section == "project") {
            sawProject = true;
            while (!c.atClose()) plan->projections.push_back(er.read(c));
            if (plan->projections.empty()) malformed("empty (project)");
        } else if (section == "aggregate") {
            plan->hasAgg = true;
            while (!c.atClose()) {
                const std::string clause = c.enterNamed();
                if (clause == "step") {
                    const std::string s = c.atom();
                    if (s == "single") plan->step = Step::SINGLE;
                    else if (s == "partial") plan->step = Step::PARTIAL;
                    else if (s == "final") plan->step = Step::FINAL;
                    else if (s == "intermediate") plan->step = Step::INTERMEDIATE;
                    else malformed("step '" + s + "'");
                } else if (clause == "group") {
                    while (!c.atClose()) plan->groupChannels.push_back((int)c.integer("group channel"));
                } else if (clause == "expected-groups") {
                    plan->expectedGroups = std::max<int64_t>(1, c.integer("expected-groups"));
                } else if (clause == "agg") {
                    plan->aggs.push_back(readAgg(c));
                } else {
                    malformed("aggregate clause '" + clause + "'");
                }
                c.leave();
            }
        } else if (section == "options") {
            while (!c.atClose()) {
                const std::string key = c.enterNamed();
                const int64_t v = c.integer(key.c_str());
                if (key == "max-partial-groups") plan->maxPartialGroups = v;
                else if (key == "max-output-rows") plan->maxOutputRows = v;
                else if (key == "merge-min-rows") plan->mergeMinRows = v;
                else if (key == "merge-min-bytes") plan->mergeMinBytes = v;
                else if (key == "merge-max-bytes") plan->mergeMaxBytes = v;
                else malformed("option '" + key + "'");
                c.leave();
            }
        } else {
            malformed("section '" + section + "'");
        }
        c.leave();
    }
    c.leave();
    if (!c.atEnd()) malformed("text after the plan");
    if (!sawProject) {  // no projection: the page goes on as it is
        for (size_t ch = 0; ch < plan->input.size(); ch++) {
            auto e = std::make_unique<Expr>();
            e->kind = ExprKind::INPUT;
            e->channel = (int)ch;
            e->type = plan->input[ch];
            plan->projections.push_back(std::move(e));
        }
    }
    // channel references of the aggregation against the projected page
    const std::vector<Type> proj = plan->projectedTypes();
    auto inRange = [&](int ch) { return ch >= 0 && ch < (int)proj.size(); };
    for (int g : plan->groupChannels)
        if (!inRange(g)) malformed("group channel " + std::to_string(g));
    if (plan->hasAgg && plan->aggs.empty() && plan->groupChannels.empty()) malformed("aggregate without groups or functions");
    for (auto& a : plan->aggs) {
        if (inputIsIntermediate(plan->step)) {
            if (!a.hasArg) malformed("FINAL-step aggregate without its first state channel");
            const std::vector<Type> st = Plan::intermediateTypes(a);
            for (size_t j = 0; j < st.size(); j++) {
                const int ch = a.arg + (int)j;
                if (!inRange(ch) || proj[(size_t)ch] != st[j]) malformed("state channel " + std::to_string(ch) + " of a FINAL-step aggregate");
            }
        } else {
            if (a.hasArg && !inRange(a.arg)) malformed("aggregate channel " + std::to_string(a.arg));
            if (a.hasArg && a.fn == AggFn::COUNT) a.inType = proj[(size_t)a.arg];
            if (a.hasArg && a.fn != AggFn::COUNT && proj[(size_t)a.arg] != a.inType) malformed("aggregate channel " + std::to_string(a.arg) + " is " + proj[(size_t)a.arg].str());
            if (a.mask >= 0 && (!inRange(a.mask) || proj[(size_t)a.mask].kind != TypeKind::BOOLEAN)) malformed("mask channel");
        }
    }
    // MergePages.java:100-105 argument checks
    if (plan->mergeMinRows < 0 || plan->mergeMinBytes < 0 || plan->mergeMaxBytes <= 0 || plan->mergeMaxBytes < plan->mergeMinBytes) malformed("merge thresholds");
    if ((plan->mergeMinRows || plan->mergeMinBytes) && plan->hasAgg) malformed("merge thresholds on an aggregation plan");
    return plan;
}

}  // namespace oracle
