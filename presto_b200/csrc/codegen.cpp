// codegen.cpp — turns a validated Plan plus the encodings of the incoming page into CUDA source:
// a traits struct `P` that the kernel templates of device/pgpu_device.cuh are instantiated with.
// Reference semantics followed, node by node:
//   AND/OR        main/sql/gen/AndCodeGenerator.java, OrCodeGenerator.java (short circuit + SQL three-valued logic)
//   BETWEEN       main/sql/gen/BetweenCodeGenerator.java (value >= min AND value <= max)
//   IN            main/sql/gen/InCodeGenerator.java (NULL if no match and any NULL operand)
//   IF/COALESCE   main/sql/gen/IfCodeGenerator.java, CoalesceCodeGenerator.java
//   calls         null-on-null convention of main/sql/gen/BytecodeUtils.java generateInvocation; operator bodies from
//                 main/type/{Bigint,Integer,Double,Decimal}Operators.java
//   filter        NULL => not selected, main/sql/gen/PageFunctionCompiler.java:542
//   accumulators  main/operator/aggregation/{DoubleSum,LongSum,Count,CountColumn,Average,DecimalSum,DecimalAverage}*.java,
//                 AbstractMinMaxAggregationFunction.java:274-306 + main/util/MinMaxCompare.java
#include "codegen.h"

#include <algorithm>
#include <cctype>
#include <cstdlib>
#include <map>
#include <set>
#include <sstream>

#include "../../include/prestogpu.h"

namespace pgpu {
namespace {

// defaults of the high-cardinality knobs (PGPU_TUNE_RUNS / PGPU_TUNE_LOCB override them: bench sweeps, A/B parity tests).
// Measured on GROUP BY l_orderkey at sf100 (profiles/r01g_summary.md): run compaction 28.8 -> 24.5 ms per pass, locality-aware
// placement 28.8 -> 70 ms (sparse keys packed side by side turn linear probing into long chains). Run compaction is the
// default for the steps that read raw rows (SINGLE, PARTIAL: 23.2 -> 19.4 ms for the PARTIAL pass of the 2-GPU run); FINAL /
// INTERMEDIATE inputs are partial states, one row per key and source, where there are no runs to fold.
constexpr int kDefaultRunsSingleStep = 1;
constexpr int kDefaultRunsPartialStep = 1;
constexpr int kDefaultRunsOtherSteps = 0;
constexpr int kDefaultLocBits = 0;
int envInt(const char* name, int dflt) {
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}

[[noreturn]] void unsupported(const std::string& m) { throw PlanError(PGPU_NOT_SUPPORTED, "not supported on the GPU path: " + m); }

struct Val {
    std::string v;   // value expression (C identifier or literal)
    std::string n;   // null flag expression; empty = never null
    std::string hi;  // second word for 16-byte var-width cells
    Type type;
    bool nullable() const { return !n.empty(); }
};

std::string ctype(const Type& t) {
    switch (t.kind) {
        case TypeKind::BOOLEAN: return "bool";
        case TypeKind::INTEGER:
        case TypeKind::DATE: return "i32";
        case TypeKind::BIGINT: return "i64";
        case TypeKind::DOUBLE: return "double";
        case TypeKind::DECIMAL: return t.precision <= 18 ? "i64" : "i128";
        default: return "u64";
    }
}

// storage element type of a flat block
std::string stype(const Type& t) {
    switch (t.kind) {
        case TypeKind::BOOLEAN: return "u8";
        case TypeKind::INTEGER:
        case TypeKind::DATE: return "i32";
        case TypeKind::BIGINT: return "i64";
        case TypeKind::DOUBLE: return "double";
        case TypeKind::DECIMAL: return t.precision <= 18 ? "i64" : "raw128";
        default: return "u8";
    }
}

std::string i128Literal(i128 v) {
    bool neg = v < 0;
    u128 m = neg ? (u128)0 - (u128)v : (u128)v;
    std::ostringstream o;
    o << "(" << (neg ? "-" : "") << "(i128)(((u128)0x" << std::hex << (uint64_t)(m >> 64) << "ull << 64) | 0x" << (uint64_t)m << "ull))";
    return o.str();
}

std::string i64Literal(int64_t v) {
    if (v == INT64_MIN) return "((i64)0x8000000000000000ull)";
    return "((i64)" + std::to_string(v) + "ll)";
}

bool canRaise(const Expr& e) {
    switch (e.kind) {
        case ExprKind::INPUT:
        case ExprKind::CONST: return false;
        case ExprKind::CALL:
            switch (e.op) {
                case Op::ADD:
                case Op::SUBTRACT:
                case Op::MULTIPLY:
                case Op::NEGATE:
                    if (e.type.kind != TypeKind::DOUBLE && !(e.type.isShortDecimal())) return true;
                    break;
                case Op::DIVIDE:
                case Op::MODULUS:
                    if (e.type.kind != TypeKind::DOUBLE) return true;
                    break;
                case Op::CAST:
                    if (e.type != e.args[0]->type && !(e.type.kind == TypeKind::DOUBLE) && !(e.type.kind == TypeKind::BIGINT && e.args[0]->type.kind == TypeKind::INTEGER)) return true;
                    break;
                default: break;
            }
            break;
        default: break;
    }
    for (auto& a : e.args)
        if (canRaise(*a)) return true;
    return false;
}

void collectChannels(const Expr& e, std::set<int>& s) {
    if (e.kind == ExprKind::INPUT) s.insert(e.channel);
    for (auto& a : e.args) collectChannels(*a, s);
}

// Can the value be NULL for some page? (inputs are assumed nullable: the accumulator/key LAYOUT must not depend on
// which pages happen to carry null vectors, only the generated code does.)
bool mayBeNull(const Expr& e) {
    switch (e.kind) {
        case ExprKind::INPUT: return true;
        case ExprKind::CONST: return e.isNull;
        case ExprKind::IS_NULL: return false;
        case ExprKind::IF: return mayBeNull(*e.args[1]) || mayBeNull(*e.args[2]);
        case ExprKind::COALESCE:
            for (auto& a : e.args)
                if (!mayBeNull(*a)) return false;
            return true;
        default:
            for (auto& a : e.args)
                if (mayBeNull(*a)) return true;
            return false;
    }
}

struct Cell {
    std::string kind;  // cnt, cntw (wide), sumd, sumi, sumdec, max, min
    int word = 0;      // first word
    int nwords = 1;
    std::vector<std::string> contrib;  // per word: expression of the u64 contribution of the current row
};

struct Gen {
    const Plan& plan;
    const std::vector<ColSig>& sig;
    std::ostringstream out;   // current code sink
    int tmpId = 0;
    int indent = 2;
    // staged mode: tile loads come from a shared-memory stage `st` (thread's first row = trow) instead of global memory
    bool stagedLoads = false;
    // staged kernel: dictionaries of at most kDcCap entries are decoded once per block into shared memory (what
    // DictionaryAwarePageProjection does once per dictionary, project/DictionaryAwarePageProjection.java:125-273):
    // channel -> cache slot. Layout: [int ok[16]] then per slot kDcCap x 8-byte values + kDcCap null bytes.
    std::map<int, int> dcSlot;
    static constexpr int kDcCap = 256, kDcHeader = 64, kDcSlotBytes = kDcCap * 8 + kDcCap, kDcMaxSlots = 8;
    bool dcEligible(const Type& t) const {
        if (t.kind == TypeKind::VARCHAR) return t.precision >= 1 && t.precision <= 7;  // one packed word, length in byte 0
        if (t.kind == TypeKind::BOOLEAN || t.isVarWidth() || t.isLongDecimal()) return false;
        return true;
    }
    int SR = 2;  // rows per thread of the staged variant (the direct variants use 4)
    struct StageOff { int values = -1, nulls = -1, offsets = -1, vbytes = -1, vbase = -1, vcap = 0; };
    std::map<int, StageOff> soff;

    Gen(const Plan& p, const std::vector<ColSig>& s) : plan(p), sig(s) {}

    std::string tmp(const char* p) { return std::string(p) + std::to_string(tmpId++); }
    void line(const std::string& s) { out << std::string(indent * 4, ' ') << s << "\n"; }

    // ------------------------------------------------------------------ column loads
    void emitTileLoads(const std::set<int>& chans) {
        for (int c : chans) {
            const Type& t = plan.input[c];
            const ColSig& cs = sig[c];
            std::string C = std::to_string(c);
            if (cs.encoding == PGPU_ENC_RLE) continue;  // loaded in the row prologue (uniform)
            if (stagedLoads) {
                const StageOff& so = soff[c];
                auto at = [](int off) { return "st + " + std::to_string(off); };
                const std::string R = std::to_string(SR);
                if (cs.encoding == PGPU_ENC_DICTIONARY) {
                    line("i32 id" + C + "[" + R + "]; lds_tile<i32, " + R + ">(" + at(so.values) + ", trow, id" + C + ");");
                    continue;
                }
                if (t.isVarWidth()) {
                    line("i32 o" + C + "[" + R + " + 1]; { i32 tq[" + R + "]; lds_tile<i32, " + R + ">(" + at(so.offsets) + ", trow, tq); for (int q = 0; q < " + R + "; q++) o" + C + "[q] = tq[q]; o" + C + "[" + R +
                         "] = *reinterpret_cast<const i32*>(" + at(so.offsets) + " + (trow + " + R + ") * 4); }");
                    line("const i32 vb" + C + " = *reinterpret_cast<const i32*>(" + at(so.vbase) + ");");
                    if (false && t.kind == TypeKind::VARCHAR && t.precision <= 4)
                        line("u64 wl" + C + ", wh" + C + "; lds_byte_window<" + std::to_string(SR * t.precision) + ">(" + at(so.vbytes) + ", o" + C + "[0] - vb" + C + ", o" + C + "[" + R + "] - o" + C + "[0], wl" + C + ", wh" +
                             C + ", err);");
                } else {
                    std::string st = stype(t);
                    line(st + " c" + C + "[" + R + "]; lds_tile<" + st + ", " + R + ">(" + at(so.values) + ", trow, c" + C + ");");
                }
                if (cs.nullable) line("u8 n" + C + "[" + R + "]; lds_tile<u8, " + R + ">(" + at(so.nulls) + ", trow, n" + C + ");");
                continue;
            }
            if (cs.encoding == PGPU_ENC_DICTIONARY) {
                line("i32 id" + C + "[4]; load_tile4<i32>(a.c[" + C + "].ids, base, nrows, full, id" + C + ");");
                continue;
            }
            if (cs.encoding == PGPU_ENC_VARCELL) {
                line("raw128 q" + C + "[4]; for (int q = 0; q < 4; q++) { raw128 e; e.lo = 0; e.hi = 0; if (base + q < nrows) e = ((const raw128*)a.c[" + C + "].values)[base + q]; q" + C + "[q] = e; }");
            } else if (t.isVarWidth()) {
                line("i32 o" + C + "[5];");
                line("if (full) { i32 t4[4]; load_tile4<i32>(a.c[" + C + "].offsets, base, nrows, true, t4); o" + C + "[0] = t4[0]; o" + C + "[1] = t4[1]; o" + C +
                     "[2] = t4[2]; o" + C + "[3] = t4[3]; o" + C + "[4] = a.c[" + C + "].offsets[base + 4]; }");
                line("else { for (int q = 0; q < 5; q++) { i64 ix = base + q < nrows ? base + q : nrows; o" + C + "[q] = a.c[" + C + "].offsets[ix]; } }");
                if (t.kind == TypeKind::VARCHAR && t.precision <= 4)  // the 4 rows' payload is one window of <= 16 bytes
                    line("u64 wl" + C + ", wh" + C + "; load_byte_window<" + std::to_string(4 * t.precision) + ">((const u8*)a.c[" + C + "].values, o" + C + "[0], o" + C + "[4] - o" + C + "[0], wl" + C + ", wh" + C + ", err);");
            } else {
                std::string st = stype(t);
                line(st + " c" + C + "[4]; load_tile4<" + st + ">((const " + st + "*)a.c[" + C + "].values, base, nrows, full, c" + C + ");");
            }
            if (cs.nullable) line("u8 n" + C + "[4]; load_tile4<u8>(a.c[" + C + "].nulls, base, nrows, full, n" + C + ");");
        }
    }

    // per-row temps x<c> (value), y<c> (2nd word of var-width), z<c> (null flag)
    void emitRowPrologue(const std::set<int>& chans) {
        for (int c : chans) {
            const Type& t = plan.input[c];
            const ColSig& cs = sig[c];
            std::string C = std::to_string(c);
            std::string ct = ctype(t), st = stype(t);
            auto conv = [&](const std::string& raw) -> std::string {
                if (t.kind == TypeKind::BOOLEAN) return "(" + raw + " != 0)";
                if (t.isLongDecimal()) return "dec_from_raw(" + raw + ")";
                return raw;
            };
            if (cs.encoding == PGPU_ENC_VARCELL) {
                line("const u64 x" + C + " = q" + C + "[r].lo, y" + C + " = q" + C + "[r].hi;");
                if (cs.nullable) line("const bool z" + C + " = n" + C + "[r] != 0;");
            } else if (cs.encoding == PGPU_ENC_PLAIN) {
                if (t.isVarWidth()) {
                    int n = t.kind == TypeKind::VARCHAR ? t.precision : 0;
                    if (t.kind == TypeKind::VARCHAR && n <= 4 && stagedLoads) {
                        line("u64 x" + C + ", y" + C + " = 0; lds_varchar<" + std::to_string(n) + ">(st + " + std::to_string(soff[c].vbytes) + ", o" + C + "[r] - vb" + C + ", o" + C + "[r + 1] - o" + C + "[r], x" + C + ", err);");
                    } else if (t.kind == TypeKind::VARCHAR && n <= 4) {
                        line("u64 x" + C + ", y" + C + " = 0; window_varchar<" + std::to_string(n) + ", " + std::to_string((stagedLoads ? SR : 4) * n) + ">(wl" + C + ", wh" + C + ", o" + C + "[r] - o" + C + "[0], o" + C + "[r + 1] - o" + C + "[r], x" + C + ", err);");
                    } else if (t.kind == TypeKind::VARCHAR && stagedLoads) {
                        line("u64 x" + C + ", y" + C + "; pack_varchar<" + std::to_string(n) + ">((const u8*)(st + " + std::to_string(soff[c].vbytes) + "), o" + C + "[r] - vb" + C + ", o" + C + "[r + 1] - o" + C + "[r], x" + C + ", y" + C + ", err);");
                    } else if (t.kind == TypeKind::VARCHAR) {
                        line("u64 x" + C + ", y" + C + "; pack_varchar<" + std::to_string(n) + ">((const u8*)a.c[" + C + "].values, o" + C + "[r], o" + C + "[r + 1] - o" + C + "[r], x" + C + ", y" + C + ", err);");
                    } else {
                        line("const u8* x" + C + " = (const u8*)a.c[" + C + "].values + o" + C + "[r];");
                    }
                } else {
                    line("const " + ct + " x" + C + " = " + conv("c" + C + "[r]") + ";");
                }
                if (cs.nullable) line("const bool z" + C + " = n" + C + "[r] != 0;");
            } else {
                std::string idx = cs.encoding == PGPU_ENC_DICTIONARY ? "id" + C + "[r]" : "0";
                if (stagedLoads && cs.encoding == PGPU_ENC_DICTIONARY && dcEligible(t) && (dcSlot.count(c) || (int)dcSlot.size() < kDcMaxSlots)) {
                    if (!dcSlot.count(c)) {
                        const int k = (int)dcSlot.size();
                        dcSlot[c] = k;
                    }
                    const int k = dcSlot[c];
                    const std::string dv = "((const u64*)(dc + " + std::to_string(kDcHeader + k * kDcSlotBytes) + "))[" + idx + "]";
                    const std::string dn = "(dc + " + std::to_string(kDcHeader + k * kDcSlotBytes + kDcCap * 8) + ")[" + idx + "]";
                    if (t.kind == TypeKind::VARCHAR) {
                        // length byte 0xff marks an entry longer than the type allows: only rows that use it raise
                        line("u64 x" + C + " = " + dv + ", y" + C + " = 0; if ((x" + C + " & 0xffull) == 0xffull) { err |= PGPU_ERR_VARCHAR_LEN; x" + C + " = 0; }");
                    } else {
                        line("const " + ct + " x" + C + " = " + (t.kind == TypeKind::DOUBLE ? "u2d(" + dv + ")" : "(" + ct + ")" + dv) + ";");
                    }
                    if (cs.nullable) line("const bool z" + C + " = " + dn + " != 0;");
                    continue;
                }
                if (t.isVarWidth()) {
                    if (t.kind == TypeKind::VARCHAR) {
                        line("u64 x" + C + ", y" + C + "; { const i32 q0 = a.c[" + C + "].doffsets[" + idx + "], q1 = a.c[" + C + "].doffsets[" + idx + " + 1]; pack_varchar<" +
                             std::to_string(t.precision) + ">((const u8*)a.c[" + C + "].dvalues, q0, q1 - q0, x" + C + ", y" + C + ", err); }");
                    } else {
                        line("const u8* x" + C + " = (const u8*)a.c[" + C + "].dvalues + a.c[" + C + "].doffsets[" + idx + "];");
                    }
                } else {
                    line("const " + ct + " x" + C + " = " + conv("((const " + st + "*)a.c[" + C + "].dvalues)[" + idx + "]") + ";");
                }
                if (cs.nullable) line("const bool z" + C + " = a.c[" + C + "].dnulls[" + idx + "] != 0;");
            }
        }
    }

    Val colVal(int c) {
        Val v;
        std::string C = std::to_string(c);
        v.type = plan.input[c];
        v.v = "x" + C;
        if (v.type.kind == TypeKind::VARCHAR) v.hi = "y" + C;
        if (sig[c].nullable) v.n = "z" + C;
        return v;
    }

    // ------------------------------------------------------------------ expressions
    Val constant(const Expr& e) {
        Val v;
        v.type = e.type;
        if (e.isNull) {
            v.n = "true";
            v.v = e.type.kind == TypeKind::DOUBLE ? "0.0" : (e.type.kind == TypeKind::BOOLEAN ? "false" : "(" + ctype(e.type) + ")0");
            return v;
        }
        switch (e.type.kind) {
            case TypeKind::BOOLEAN: v.v = e.i64 ? "true" : "false"; break;
            case TypeKind::INTEGER:
            case TypeKind::DATE: v.v = "((i32)" + std::to_string(e.i64) + ")"; break;
            case TypeKind::BIGINT: v.v = i64Literal(e.i64); break;
            case TypeKind::DOUBLE: {
                std::ostringstream o;
                o << "u2d(0x" << std::hex << (uint64_t)e.i64 << "ull)";
                v.v = o.str();
                break;
            }
            case TypeKind::DECIMAL: v.v = e.type.isShortDecimal() ? i64Literal((int64_t)e.dec) : i128Literal(e.dec); break;
            default: unsupported("constant of type " + e.type.str());
        }
        return v;
    }

    static std::string nullOr(const std::string& a, const std::string& b) {
        if (a.empty()) return b;
        if (b.empty()) return a;
        return "(" + a + " || " + b + ")";
    }

    Val compare(Op op, const Val& a, const Val& b) {
        Val r;
        r.type = Type::boolean();
        const char* o = "==";
        switch (op) {
            case Op::EQ: o = "=="; break;
            case Op::NE: o = "!="; break;
            case Op::LT: o = "<"; break;
            case Op::LE: o = "<="; break;
            case Op::GT: o = ">"; break;
            case Op::GE: o = ">="; break;
            default: break;
        }
        std::string t = tmp("b");
        line("const bool " + t + " = (" + a.v + " " + o + " " + b.v + ");");
        r.v = t;
        r.n = nullOr(a.n, b.n);
        return r;
    }

    // three-valued AND / OR over already evaluated operands
    Val logic(bool isAnd, const std::vector<Val>& ops) {
        Val r;
        r.type = Type::boolean();
        bool anyNullable = false;
        for (auto& o : ops) anyNullable |= o.nullable();
        std::string t = tmp("b");
        if (!anyNullable) {
            std::string e;
            for (size_t i = 0; i < ops.size(); i++) e += (i ? (isAnd ? " && " : " || ") : "") + ops[i].v;
            line("const bool " + t + " = (" + e + ");");
            r.v = t;
            return r;
        }
        // decisive = some operand is definitely false (AND) / true (OR)
        std::string dec, anyn;
        for (size_t i = 0; i < ops.size(); i++) {
            std::string d = isAnd ? "!" + ops[i].v : ops[i].v;
            if (ops[i].nullable()) d = "(!" + ops[i].n + " && " + d + ")";
            dec += (i ? " || " : "") + d;
            if (ops[i].nullable()) anyn += (anyn.empty() ? "" : " || ") + ops[i].n;
        }
        std::string td = tmp("b"), tn = tmp("b");
        line("const bool " + td + " = (" + dec + ");");
        line("const bool " + tn + " = !" + td + " && (" + anyn + ");");
        line("const bool " + t + " = " + (isAnd ? "!" + td + " && !" + tn : td) + ";");
        r.v = t;
        r.n = tn;
        return r;
    }

    Val genLogic(const Expr& e) {
        bool isAnd = e.kind == ExprKind::AND;
        bool guard = false;
        for (size_t i = 1; i < e.args.size(); i++) guard |= canRaise(*e.args[i]);
        if (!guard) {
            std::vector<Val> ops;
            for (auto& a : e.args) ops.push_back(gen(*a));
            return logic(isAnd, ops);
        }
        // short circuit: later operands are only evaluated while no decisive value was seen
        std::string td = tmp("b"), tn = tmp("b");
        line("bool " + td + " = false, " + tn + " = false;");
        int opened = 0;
        for (size_t i = 0; i < e.args.size(); i++) {
            if (i > 0) {
                line("if (!" + td + ") {");
                indent++;
                opened++;
            }
            Val o = gen(*e.args[i]);
            std::string d = isAnd ? "!" + o.v : o.v;
            if (o.nullable()) {
                line("if (" + o.n + ") " + tn + " = true; else if (" + d + ") " + td + " = true;");
            } else {
                line("if (" + d + ") " + td + " = true;");
            }
        }
        while (opened--) {
            indent--;
            line("}");
        }
        Val r;
        r.type = Type::boolean();
        std::string t = tmp("b"), t2 = tmp("b");
        line("const bool " + t2 + " = !" + td + " && " + tn + ";");
        line("const bool " + t + " = " + (isAnd ? "!" + td + " && !" + tn : td) + ";");
        r.v = t;
        r.n = t2;
        return r;
    }

    Val genIf(const Expr& e) {
        Val c = gen(*e.args[0]);
        std::string cond = c.nullable() ? "(" + c.v + " && !" + c.n + ")" : c.v;
        std::string ct = ctype(e.type);
        std::string t = tmp("v"), tn = tmp("b");
        line(ct + " " + t + " = " + (e.type.kind == TypeKind::DOUBLE ? "0.0" : "(" + ct + ")0") + "; bool " + tn + " = false;");
        line("if (" + cond + ") {");
        indent++;
        Val a = gen(*e.args[1]);
        line(t + " = " + a.v + ";" + (a.nullable() ? " " + tn + " = " + a.n + ";" : ""));
        indent--;
        line("} else {");
        indent++;
        Val b = gen(*e.args[2]);
        line(t + " = " + b.v + ";" + (b.nullable() ? " " + tn + " = " + b.n + ";" : ""));
        indent--;
        line("}");
        Val r;
        r.type = e.type;
        r.v = t;
        if (a.nullable() || b.nullable()) r.n = tn;
        return r;
    }

    Val genCoalesce(const Expr& e) {
        std::string ct = ctype(e.type);
        std::string t = tmp("v"), tn = tmp("b");
        line(ct + " " + t + " = " + (e.type.kind == TypeKind::DOUBLE ? "0.0" : "(" + ct + ")0") + "; bool " + tn + " = true;");
        int opened = 0;
        bool certain = false;
        for (size_t i = 0; i < e.args.size() && !certain; i++) {
            if (i > 0) {
                line("if (" + tn + ") {");
                indent++;
                opened++;
            }
            Val a = gen(*e.args[i]);
            if (a.nullable()) {
                line("if (!" + a.n + ") { " + t + " = " + a.v + "; " + tn + " = false; }");
            } else {
                line(t + " = " + a.v + "; " + tn + " = false;");
                certain = true;
            }
        }
        while (opened--) {
            indent--;
            line("}");
        }
        Val r;
        r.type = e.type;
        r.v = t;
        if (!certain) r.n = tn;
        return r;
    }

    Val genIn(const Expr& e) {
        Val v = gen(*e.args[0]);
        std::string match, anyNull;
        for (size_t i = 1; i < e.args.size(); i++) {
            Val c = constant(*e.args[i]);
            if (e.args[i]->isNull) {
                anyNull = "true";
                continue;
            }
            match += (match.empty() ? "" : " || ") + std::string("(") + v.v + " == " + c.v + ")";
        }
        if (match.empty()) match = "false";
        std::string t = tmp("b");
        line("const bool " + t + " = (" + match + ");");
        Val r;
        r.type = Type::boolean();
        r.v = t;
        std::string n = v.n;
        if (!anyNull.empty()) {
            std::string tn = tmp("b");
            line("const bool " + tn + " = " + (v.nullable() ? "(" + v.n + " || !" + t + ")" : "!" + t) + ";");
            n = tn;
        }
        r.n = n;
        return r;
    }

    Val toI128(const Val& a) {
        if (a.type.isLongDecimal()) return a;
        Val r = a;
        r.v = "((i128)" + a.v + ")";
        return r;
    }

    Val genDecimalArith(const Expr& e, const Val& a, const Val& b) {
        const Type &ta = a.type, &tb = b.type, &tr = e.type;
        Val r;
        r.type = tr;
        r.n = nullOr(a.n, b.n);
        std::string t = tmp("v");
        if (e.op == Op::ADD || e.op == Op::SUBTRACT) {
            int s = tr.scale;
            const char* o = e.op == Op::ADD ? "+" : "-";
            if (tr.isShortDecimal()) {  // DecimalOperators.addShortShortShort :103-106, subtractShortShortShort :208-211
                line("const i64 " + t + " = " + a.v + " * " + i64Literal((int64_t)pow10_i128(s - ta.scale)) + " " + o + " " + b.v + " * " + i64Literal((int64_t)pow10_i128(s - tb.scale)) + ";");
            } else {  // internalAddLongLongLong :151-172: rescale, add, throwIfOverflows
                std::string la = toI128(a).v, lb = toI128(b).v;
                if (s > ta.scale) la = "dec_upscale(" + la + ", " + i128Literal(pow10_i128(s - ta.scale)) + ", err)";
                if (s > tb.scale) lb = "dec_upscale(" + lb + ", " + i128Literal(pow10_i128(s - tb.scale)) + ", err)";
                line("const i128 " + t + " = " + la + " " + o + " " + lb + "; if (dec_overflows(" + t + ")) err |= PGPU_ERR_OVERFLOW;");
            }
        } else if (e.op == Op::MULTIPLY) {
            if (tr.isShortDecimal()) line("const i64 " + t + " = " + a.v + " * " + b.v + ";");  // multiplyShortShortShort :281-284
            else line("const i128 " + t + " = dec_mul(" + toI128(a).v + ", " + toI128(b).v + ", err);");
        } else {
            unsupported(std::string("decimal ") + opName(e.op));
        }
        r.v = t;
        return r;
    }

    // Faults of code that the reference would not have run. BytecodeUtils.generateFullInvocation (main/sql/gen/BytecodeUtils.java:
    // 244-310) evaluates a call's arguments left to right and jumps past the rest of the call as soon as one is NULL
    // (ifWasNullPopAndGoto after every NEVER_NULL argument): a later argument is not evaluated, the operator is not invoked, so
    // neither can raise. The straight-line code here evaluates everything (values under NULL positions are arbitrary), so the
    // error bits of the code emitted since `mark` go to a private word that only counts when `nullExpr` is false.
    void guardFaults(size_t mark, const std::string& nullExpr) {
        if (nullExpr.empty()) return;
        std::string all = out.str();
        std::string tail = all.substr(mark);
        std::string renamed;
        const std::string name = tmp("e");
        bool any = false;
        for (size_t i = 0; i < tail.size();) {
            const bool startsWord = i == 0 || !(isalnum((unsigned char)tail[i - 1]) || tail[i - 1] == '_');
            if (startsWord && tail.compare(i, 3, "err") == 0 && (i + 3 >= tail.size() || !(isalnum((unsigned char)tail[i + 3]) || tail[i + 3] == '_'))) {
                renamed += name;
                i += 3;
                any = true;
            } else {
                renamed += tail[i++];
            }
        }
        if (!any) return;
        const std::string pad(indent * 4, ' ');
        out.str(all.substr(0, mark) + pad + "u32 " + name + " = 0;\n" + renamed + pad + "if (!(" + nullExpr + ")) err |= " + name + ";\n");
        out.seekp(0, std::ios::end);
    }

    Val genCall(const Expr& e) {
        std::vector<Val> a;
        std::string earlierNull;  // some earlier argument of this call is NULL
        for (auto& x : e.args) {
            const size_t mark = (size_t)out.tellp();
            a.push_back(gen(*x));
            guardFaults(mark, earlierNull);
            earlierNull = nullOr(earlierNull, a.back().n);
        }
        const size_t mark = (size_t)out.tellp();
        Val r = genCallOp(e, a);
        guardFaults(mark, earlierNull);
        return r;
    }

    Val genCallOp(const Expr& e, std::vector<Val>& a) {
        Val r;
        r.type = e.type;
        for (auto& x : a) r.n = nullOr(r.n, x.n);
        std::string t = tmp("v");
        std::string ct = ctype(e.type);
        switch (e.op) {
            case Op::EQ:
            case Op::NE:
            case Op::LT:
            case Op::LE:
            case Op::GT:
            case Op::GE: return compare(e.op, a[0], a[1]);
            case Op::NOT:
                line("const bool " + t + " = !" + a[0].v + ";");
                r.v = t;
                return r;
            case Op::NEGATE:
                if (e.type.kind == TypeKind::DOUBLE || e.type.kind == TypeKind::DECIMAL) line("const " + ct + " " + t + " = -" + a[0].v + ";");
                else if (e.type.kind == TypeKind::BIGINT) line("const i64 " + t + " = neg_i64_checked(" + a[0].v + ", err);");
                else line("const i32 " + t + " = narrow_i32_checked(-(i64)" + a[0].v + ", err);");
                r.v = t;
                return r;
            case Op::ADD:
            case Op::SUBTRACT:
            case Op::MULTIPLY:
            case Op::DIVIDE:
            case Op::MODULUS: {
                if (e.type.kind == TypeKind::DECIMAL) return genDecimalArith(e, a[0], a[1]);
                const char* o = e.op == Op::ADD ? "+" : e.op == Op::SUBTRACT ? "-" : e.op == Op::MULTIPLY ? "*" : "/";
                if (e.type.kind == TypeKind::DOUBLE) {
                    if (e.op == Op::MODULUS) line("const double " + t + " = fmod(" + a[0].v + ", " + a[1].v + ");");
                    else line("const double " + t + " = " + a[0].v + " " + o + " " + a[1].v + ";");
                } else if (e.type.kind == TypeKind::BIGINT) {
                    const char* f = e.op == Op::ADD ? "add" : e.op == Op::SUBTRACT ? "sub" : e.op == Op::MULTIPLY ? "mul" : e.op == Op::DIVIDE ? "div" : "mod";
                    line("const i64 " + t + " = " + f + "_i64_checked(" + a[0].v + ", " + a[1].v + ", err);");
                } else {  // INTEGER (main/type/IntegerOperators.java:45-103)
                    if (e.op == Op::DIVIDE) line("const i32 " + t + " = div_i32_checked(" + a[0].v + ", " + a[1].v + ", err);");
                    else if (e.op == Op::MODULUS) line("const i32 " + t + " = mod_i32_checked(" + a[0].v + ", " + a[1].v + ", err);");
                    else line("const i32 " + t + " = narrow_i32_checked((i64)" + a[0].v + " " + o + " (i64)" + a[1].v + ", err);");
                }
                r.v = t;
                return r;
            }
            case Op::CAST: {
                const Type &f = a[0].type, &to = e.type;
                if (f == to) return a[0];
                auto k = [](const Type& x) { return x.kind; };
                if (k(f) == TypeKind::INTEGER && k(to) == TypeKind::BIGINT) line("const i64 " + t + " = (i64)" + a[0].v + ";");
                else if ((k(f) == TypeKind::INTEGER || k(f) == TypeKind::BIGINT) && k(to) == TypeKind::DOUBLE) line("const double " + t + " = (double)" + a[0].v + ";");
                else if (k(f) == TypeKind::BIGINT && k(to) == TypeKind::INTEGER) line("const i32 " + t + " = narrow_i32_checked(" + a[0].v + ", err);");
                else if (k(f) == TypeKind::DOUBLE && k(to) == TypeKind::BIGINT) line("const i64 " + t + " = cast_f64_i64(" + a[0].v + ", err);");
                else if (k(f) == TypeKind::BOOLEAN && (k(to) == TypeKind::BIGINT || k(to) == TypeKind::INTEGER)) line("const " + ct + " " + t + " = " + a[0].v + " ? 1 : 0;");
                else if (k(f) == TypeKind::BOOLEAN && k(to) == TypeKind::DOUBLE) line("const double " + t + " = " + a[0].v + " ? 1.0 : 0.0;");
                else if ((k(f) == TypeKind::BIGINT || k(f) == TypeKind::INTEGER) && k(to) == TypeKind::BOOLEAN) line("const bool " + t + " = " + a[0].v + " != 0;");
                else if (k(f) == TypeKind::DOUBLE && k(to) == TypeKind::BOOLEAN) line("const bool " + t + " = " + a[0].v + " != 0.0;");
                else if ((k(f) == TypeKind::BIGINT || k(f) == TypeKind::INTEGER) && k(to) == TypeKind::DECIMAL) {
                    // main/type/DecimalCasts.java bigintToShortDecimal / bigintToLongDecimal: value * 10^scale, precision check
                    if (to.isShortDecimal()) {
                        std::string lim = i64Literal((int64_t)pow10_i128(to.precision - to.scale));
                        line("const i64 " + t + " = (i64)" + a[0].v + " * " + i64Literal((int64_t)pow10_i128(to.scale)) + "; if ((i64)" + a[0].v + " >= " + lim + " || (i64)" + a[0].v + " <= -" + lim + ") err |= PGPU_ERR_OVERFLOW;");
                    } else {
                        std::string lim = i128Literal(pow10_i128(to.precision));
                        line("const i128 " + t + " = (i128)" + a[0].v + " * " + i128Literal(pow10_i128(to.scale)) + "; if (" + t + " >= " + lim + " || " + t + " <= -" + lim + ") err |= PGPU_ERR_OVERFLOW;");
                    }
                } else if (k(f) == TypeKind::DECIMAL && k(to) == TypeKind::DECIMAL) {
                    // main/type/DecimalToDecimalCasts.java: rescale (HALF_UP when narrowing) then precision check
                    std::string src = toI128(a[0]).v, w = tmp("v");
                    if (to.scale >= f.scale) line("const i128 " + w + " = " + src + " * " + i128Literal(pow10_i128(to.scale - f.scale)) + ";");
                    else line("const i128 " + w + " = dec_div_round(" + src + ", " + i128Literal(pow10_i128(f.scale - to.scale)) + ");");
                    std::string lim = i128Literal(pow10_i128(to.precision));
                    line("if (" + w + " >= " + lim + " || " + w + " <= -" + lim + ") err |= PGPU_ERR_OVERFLOW;");
                    if (f.precision - f.scale + to.scale > 38) unsupported("decimal cast whose intermediate exceeds 38 digits");
                    line("const " + ct + " " + t + " = (" + ct + ")" + w + ";");
                } else {
                    unsupported("cast " + f.str() + " -> " + to.str());
                }
                r.v = t;
                return r;
            }
        }
        unsupported("operator");
    }

    Val gen(const Expr& e) {
        switch (e.kind) {
            case ExprKind::INPUT: return colVal(e.channel);
            case ExprKind::CONST: return constant(e);
            case ExprKind::CALL: return genCall(e);
            case ExprKind::AND:
            case ExprKind::OR: return genLogic(e);
            case ExprKind::BETWEEN: {
                // BetweenCodeGenerator: value >= min AND value <= max, the value evaluated once. min is the second argument of the first
                // comparison (not evaluated when the value is NULL); max belongs to the second comparison, which AND reaches only when
                // the first is not FALSE: faults of the bounds count under those conditions only.
                Val v = gen(*e.args[0]);
                const size_t m1 = (size_t)out.tellp();
                Val lo = gen(*e.args[1]);
                guardFaults(m1, v.n);
                Val ge = compare(Op::GE, v, lo);
                const size_t m2 = (size_t)out.tellp();
                Val hi = gen(*e.args[2]);
                guardFaults(m2, nullOr(v.n, ge.nullable() ? "(!" + ge.n + " && !" + ge.v + ")" : "(!" + ge.v + ")"));
                Val le = compare(Op::LE, v, hi);
                return logic(true, {ge, le});
            }
            case ExprKind::IN: return genIn(e);
            case ExprKind::IS_NULL: {
                Val v = gen(*e.args[0]);
                Val r;
                r.type = Type::boolean();
                r.v = v.nullable() ? v.n : "false";
                return r;
            }
            case ExprKind::IF: return genIf(e);
            case ExprKind::COALESCE: return genCoalesce(e);
        }
        unsupported("expression");
    }
};

// ---------------------------------------------------------------------------------------------- key packing
struct KeyPart {
    std::string expr;  // u64 expression (already masked to nbytes)
    int nbytes;
};

struct KeyField {
    int projChannel;
    Type type;
    bool nullable;
    int nullBit = -1;      // bit in the null bytes
    int byteOffset = 0;    // of the value bytes
    int nbytes = 0;
};

int keyBytes(const Type& t) {
    switch (t.kind) {
        case TypeKind::BOOLEAN: return 1;
        case TypeKind::INTEGER:
        case TypeKind::DATE: return 4;
        case TypeKind::BIGINT:
        case TypeKind::DOUBLE: return 8;
        case TypeKind::DECIMAL: return t.precision <= 18 ? 8 : 16;
        case TypeKind::VARCHAR: return t.precision + 1;
        default: return 0;
    }
}

std::string maskOf(int nbytes) {
    if (nbytes >= 8) return "0xffffffffffffffffull";
    std::ostringstream o;
    o << "0x" << std::hex << ((1ull << (8 * nbytes)) - 1) << "ull";
    return o.str();
}

}  // namespace

GenResult generate(const Plan& plan, const std::vector<ColSig>& sig, const GenOptions& opt) {
    if (sig.size() != plan.input.size()) throw PlanError(PGPU_INVALID_ARGUMENT, "page has " + std::to_string(sig.size()) + " channels, plan expects " + std::to_string(plan.input.size()));
    GenResult res;
    res.NC = (int)plan.input.size();
    Gen g(plan, sig);
    std::vector<Type> proj = plan.projectedTypes();
    const bool finalStep = plan.hasAgg && inputIsIntermediate(plan.step);  // FINAL / INTERMEDIATE: inputs are partial states

    // which projected channels does the consumer need?
    std::set<int> neededProj;
    if (plan.hasAgg) {
        for (int c : plan.groupChannels) neededProj.insert(c);
        for (auto& a : plan.aggs) {
            if (a.hasArg) {
                int n = finalStep ? (int)Plan::intermediateTypes(a).size() : 1;
                for (int j = 0; j < n; j++) neededProj.insert(a.arg + j);
            }
            if (a.mask >= 0) neededProj.insert(a.mask);
        }
    } else {
        for (size_t c = 0; c < proj.size(); c++) neededProj.insert((int)c);
    }
    std::set<int> filterChans, restChans;
    if (plan.filter) collectChannels(*plan.filter, filterChans);
    for (int c : neededProj) collectChannels(*plan.projections[c], restChans);
    std::set<int> phase2Loads;
    for (int c : restChans)
        if (!filterChans.count(c)) phase2Loads.insert(c);
    for (int c : filterChans) res.usedChannels.push_back(c);
    for (int c : phase2Loads) res.usedChannels.push_back(c);
    for (int c : res.usedChannels) {
        const Type& t = plan.input[c];
        if (t.kind == TypeKind::VARCHAR && (t.precision == 0 || t.precision > 15))
            unsupported("reading " + t.str() + " values (GPU var-width cells hold varchar(n), n <= 15)");
        if (t.kind == TypeKind::VARBINARY && !finalStep) unsupported("varbinary input outside FINAL-step intermediate states");
    }
    bool projRaises = false;
    for (int c : neededProj) projRaises |= canRaise(*plan.projections[c]);

    std::ostringstream src;
    src << "// generated by presto_b200 codegen — plan:\n";
    {
        std::istringstream pt(plan.text);
        std::string l;
        while (std::getline(pt, l)) src << "//   " << l << "\n";
    }
    src << "#include \"pgpu_device.cuh\"\n\n";

    // ---------------------------------------------------------------- eval prologue shared by all modes
    // ---- stage layout of the staged variant: one region per used array, 128-byte aligned
    constexpr int kStagedWarps = 15;               // consumer warps (+ 1 producer warp = 512 threads, 128 registers each)
    // sub-tiles per stage: wide rows (>= 40 bytes of staged arrays per row, Q1) run fastest with two 960-row sub-tiles per
    // stage and a 2-deep ring (measured: 5.6 -> 4.1 ms per sf100 pass), narrow rows with one sub-tile and a 4-deep ring
    int rowBytes = 0;
    for (int c : res.usedChannels) {
        const Type& t = plan.input[c];
        if (sig[c].encoding == PGPU_ENC_RLE) continue;
        rowBytes += sig[c].encoding == PGPU_ENC_DICTIONARY ? 4 : t.isVarWidth() ? 4 + std::max(t.precision, 1) : t.width();
        if (sig[c].encoding == PGPU_ENC_PLAIN && sig[c].nullable) rowBytes += 1;
    }
    const int kSub = getenv("PGPU_TUNE_SUB") ? std::max(1, atoi(getenv("PGPU_TUNE_SUB"))) : (rowBytes >= 40 && rowBytes <= 100 ? 2 : 1);
    const int kTile = kStagedWarps * 32 * g.SR * kSub;  // rows per stage: every consumer thread owns SR consecutive rows of each sub-tile
    bool stagedPossible = true;
    int stageBytes = 0, tailGuard = 0;
    std::ostringstream issueLoad, issueVar, issueCopies;  // stage_issue(): global loads before the stage is free; stores into the stage; bulk copies of lane 0
    std::ostringstream preFields, preLoads, limFields, limLoads;  // StagePre / StageLim of the producer
    unsigned fixedTx = 0;
    {
        auto region = [&](int bytes) {
            int off = stageBytes;
            stageBytes += (bytes + 127) / 128 * 128;
            return off;
        };
        auto copy = [&](int off, const std::string& src, int bytes) {
            issueCopies << "            bulk_g2s(st + " << off << ", " << src << ", " << bytes << ", bar);\n";
            fixedTx += (unsigned)bytes;
        };
        for (int c : res.usedChannels) {
            const Type& t = plan.input[c];
            const ColSig& cs = sig[c];
            std::string C = std::to_string(c);
            Gen::StageOff so;
            if (cs.encoding == PGPU_ENC_RLE) continue;
            if (cs.encoding == PGPU_ENC_VARCELL) { stagedPossible = false; break; }  // device-to-device pages of partial results: short
            if (cs.encoding == PGPU_ENC_DICTIONARY) {
                so.values = region(kTile * 4);
                copy(so.values, "(const char*)a.c[" + C + "].ids + row0 * 4", kTile * 4);
            } else if (t.isVarWidth()) {
                if (t.kind != TypeKind::VARCHAR || t.precision > 4) { stagedPossible = false; break; }
                tailGuard = 3;
                so.offsets = region(kTile * 4 + 16);
                copy(so.offsets, "(const char*)a.c[" + C + "].offsets + row0 * 4", kTile * 4 + 16);
                so.vcap = (kTile * t.precision + 15) / 16 * 16 + 16;
                so.vbytes = region(so.vcap + 32);
                so.vbase = region(16);
                // byte range of the tile: 16-byte aligned body by bulk copy, the < 16 trailing bytes by the producer lanes
                // StagePre: the two bounding offsets, loaded by the producer two tiles ahead; StageLim: end of the readable bytes
                preFields << "        i32 vb0_" << C << ", vb1_" << C << ";\n";
                preLoads << "        p.vb0_" << C << " = a.c[" << C << "].offsets[row0]; p.vb1_" << C << " = a.c[" << C << "].offsets[row0 + " << kTile << "];\n";
                limFields << "        i32 vend_" << C << ";\n";
                limLoads << "        l.vend_" << C << " = a.c[" << C << "].offsets[a.rows + " << 3 << "];\n";
                // the copy end is rounded UP to 16 bytes when the over-read stays inside the byte array (always, except in
                // the last tiles of a page); otherwise rounded down and the < 16 trailing bytes are moved by hand
                issueLoad << "        const i32 vb0_" << C << " = p.vb0_" << C << ", vb1_" << C << " = p.vb1_" << C << ";\n";
                issueLoad << "        i32 s0_" << C << " = vb0_" << C << " & ~15, e1_" << C << " = (vb1_" << C << " + 15) & ~15;\n";
                issueLoad << "        const bool up_" << C << " = e1_" << C << " <= l.vend_" << C << ";\n";
                issueLoad << "        if (!up_" << C << ") e1_" << C << " = vb1_" << C << " & ~15;\n";
                issueLoad << "        if (e1_" << C << " < s0_" << C << ") e1_" << C << " = s0_" << C << ";\n";
                issueLoad << "        if (e1_" << C << " - s0_" << C << " > " << (so.vcap + 16) << ") { err |= PGPU_ERR_VARCHAR_LEN; e1_" << C << " = s0_" << C << " + " << so.vcap << "; }\n";
                issueLoad << "        const i32 tq_" << C << " = e1_" << C << " + lane; const bool tl_" << C << " = !up_" << C << " && lane < 16 && tq_" << C << " < vb1_" << C << " && tq_" << C << " - s0_" << C << " < " << (so.vcap + 16) << ";\n";
                issueLoad << "        u8 tb_" << C << " = 0; if (tl_" << C << ") tb_" << C << " = ((const u8*)a.c[" << C << "].values)[tq_" << C << "];\n";
                issueVar << "        if (tl_" << C << ") st[" << so.vbytes << " + (tq_" << C << " - s0_" << C << ")] = tb_" << C << ";\n";
                issueVar << "        if (lane == 0) *reinterpret_cast<i32*>(st + " << so.vbase << ") = s0_" << C << ";\n";
                issueVar << "        tx += (u32)(e1_" << C << " - s0_" << C << ");\n";
                issueCopies << "            if (e1_" << C << " > s0_" << C << ") bulk_g2s(st + " << so.vbytes << ", (const char*)a.c[" << C << "].values + s0_" << C << ", (u32)(e1_" << C << " - s0_" << C << "), bar);\n";
            } else {
                so.values = region(kTile * t.width());
                copy(so.values, "(const char*)a.c[" + C + "].values + row0 * " + std::to_string(t.width()), kTile * t.width());
            }
            if (cs.encoding == PGPU_ENC_PLAIN && cs.nullable) {
                so.nulls = region(kTile);
                copy(so.nulls, "(const char*)a.c[" + C + "].nulls + row0", kTile);
            }
            g.soff[c] = so;
        }
        if (res.usedChannels.empty()) stagedPossible = false;
    }

    std::ostringstream evalFilter, evalFilterEager, evalFilterStaged;  // code: loads + per-row filter -> sel[]
    for (int eager = 0; eager < 3; eager++) {
        if (eager == 2 && !stagedPossible) break;
        g.stagedLoads = eager == 2;
        g.out.str("");
        g.indent = 2;
        const std::string R = std::to_string(eager == 2 ? g.SR : 4);
        if (eager == 2) g.line("const i64 base = 0, nrows = " + R + ";  // staged tiles are always complete");
        g.line("const bool full = base + " + R + " <= nrows;");
        g.emitTileLoads(filterChans);
        if (eager) g.emitTileLoads(phase2Loads);
        g.line("bool any = false;");
        g.line("#pragma unroll");
        g.line("for (int r = 0; r < " + R + "; r++) {");
        g.indent++;
        g.line("const bool live = base + r < nrows;");
        if (plan.filter) {
            bool fr = canRaise(*plan.filter);
            if (fr) {
                g.line("bool f = false;");
                g.line("if (live) {");
                g.indent++;
            }
            g.emitRowPrologue(filterChans);
            Val f = g.gen(*plan.filter);
            std::string fe = f.nullable() ? "(" + f.v + " && !" + f.n + ")" : f.v;
            if (fr) {
                g.line("f = " + fe + ";");
                g.indent--;
                g.line("}");
                g.line("sel[r] = f;");
            } else {
                g.line("sel[r] = live && " + fe + ";");
            }
        } else {
            g.line("sel[r] = live;");
        }
        g.line("any |= sel[r];");
        g.indent--;
        g.line("}");
        g.line("if (!any) return;");
        if (!eager) g.emitTileLoads(phase2Loads);
        (eager == 2 ? evalFilterStaged : eager ? evalFilterEager : evalFilter) << g.out.str();
    }
    g.stagedLoads = false;

    auto genProjections = [&](std::map<int, Val>& pv) {
        g.emitRowPrologue(restChans);
        for (int c : neededProj) pv[c] = g.gen(*plan.projections[c]);
    };

    if (!plan.hasAgg) {
        // ============================================================ COMPACT
        res.mode = KernelMode::COMPACT;
        res.NO = (int)proj.size();
        for (auto& t : proj) {
            OutColDesc d;
            d.type = t;
            if (t.kind == TypeKind::VARCHAR) {
                d.stride = 16;
                d.varWidth = true;
                d.lenPrefixed = true;
            } else {
                d.stride = t.width();
            }
            res.out.push_back(d);
        }
        src << "struct P {\n";
        src << "    static constexpr int NC = " << res.NC << ", NO = " << res.NO << ", KW = 0, NW = 0, SLOTW = 4;\n";
        src << "    struct Proj {\n";
        for (size_t c = 0; c < proj.size(); c++) {
            src << "        " << ctype(proj[c]) << " v" << c << ";" << (proj[c].kind == TypeKind::VARCHAR ? " u64 w" + std::to_string(c) + ";" : "") << " bool n" << c << ";\n";
        }
        src << "    };\n";
        src << "    PGPU_DEV static void eval_project(const PageArgs<NC>& a, i64 base, i64 nrows, bool (&sel)[4], Proj (&pv)[4], u32& err) {\n";
        src << evalFilter.str();
        g.out.str("");
        g.indent = 2;
        g.line("#pragma unroll");
        g.line("for (int r = 0; r < 4; r++) {");
        g.indent++;
        g.line("if (sel[r]) {");
        g.indent++;
        std::map<int, Val> pv;
        genProjections(pv);
        for (size_t c = 0; c < proj.size(); c++) {
            const Val& v = pv[(int)c];
            std::string C = std::to_string(c);
            g.line("pv[r].v" + C + " = " + v.v + ";" + (proj[c].kind == TypeKind::VARCHAR ? " pv[r].w" + C + " = " + v.hi + ";" : "") + " pv[r].n" + C + " = " + (v.nullable() ? v.n : "false") + ";");
        }
        g.indent--;
        g.line("}");
        g.indent--;
        g.line("}");
        src << g.out.str();
        src << "    }\n";
        src << "    PGPU_DEV static void store_row(const OutArgs<NO>& out, i64 pos, const Proj& p) {\n";
        for (size_t c = 0; c < proj.size(); c++) {
            std::string C = std::to_string(c);
            const Type& t = proj[c];
            if (t.kind == TypeKind::VARCHAR) src << "        { raw128 q; q.lo = p.v" << C << "; q.hi = p.w" << C << "; ((raw128*)out.c[" << C << "].values)[pos] = q; }\n";
            else if (t.isLongDecimal()) src << "        ((raw128*)out.c[" << C << "].values)[pos] = dec_to_raw(p.v" << C << ");\n";
            else if (t.kind == TypeKind::BOOLEAN) src << "        ((u8*)out.c[" << C << "].values)[pos] = p.v" << C << " ? 1 : 0;\n";
            else src << "        ((" << stype(t) << "*)out.c[" << C << "].values)[pos] = p.v" << C << ";\n";
            src << "        out.c[" << C << "].nulls[pos] = p.n" << C << " ? 1 : 0;\n";
        }
        src << "    }\n";
        src << "};\n\n";
        src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_page(PageArgs<P::NC> a, OutArgs<P::NO> out, Scratch sc) { body_compact<P>(a, out, sc); }\n";
        res.source = src.str();
        return res;
    }

    // ================================================================ aggregation modes
    res.mode = plan.groupChannels.empty() ? KernelMode::AGG_GLOBAL : KernelMode::AGG_HASH;

    // The row body is generated first into a buffer because cells are discovered while generating it.
    // (its per-row column prologue exists twice: values from registers loaded out of global memory / out of a stage)
    std::string rowPrologue, rowPrologueStaged;
    g.indent = 4;
    g.out.str("");
    g.emitRowPrologue(restChans);
    rowPrologue = g.out.str();
    if (stagedPossible) {
        g.out.str("");
        g.stagedLoads = true;
        g.emitRowPrologue(restChans);
        g.stagedLoads = false;
        rowPrologueStaged = g.out.str();
    }
    g.out.str("");
    std::map<int, Val> pv;
    for (int c : neededProj) pv[c] = g.gen(*plan.projections[c]);

    // ---- keys
    std::vector<KeyField> fields;
    int nullBits = 0;
    for (int c : plan.groupChannels) {
        KeyField f;
        f.projChannel = c;
        f.type = proj[c];
        f.nullable = mayBeNull(*plan.projections[c]);
        if (f.nullable) f.nullBit = nullBits++;
        f.nbytes = keyBytes(f.type);
        fields.push_back(f);
    }
    // distinct-key bound from the key types alone (BOOLEAN, short VARCHAR): lets the host size the table once and skip
    // the per-chunk "could this chunk overflow the table" hand-shake
    {
        double dom = fields.empty() ? 0 : 1;
        for (auto& f : fields) {
            double d = 0;
            if (f.type.kind == TypeKind::BOOLEAN) d = 2;
            else if (f.type.kind == TypeKind::VARCHAR && f.type.precision <= 2) d = f.type.precision == 1 ? 257 : 65793;
            if (d == 0) { dom = 0; break; }
            dom *= d + (f.nullable ? 1 : 0);
        }
        res.keyDomain = dom;
    }
    int off = (nullBits + 7) / 8;
    for (auto& f : fields) {
        f.byteOffset = off;
        off += f.nbytes;
    }
    const int keyLen = off;
    res.KW = (keyLen + 7) / 8;
    if (res.KW > 8) unsupported("group key wider than 64 bytes");
    if (res.mode == KernelMode::AGG_HASH) {
        for (int w = 0; w < res.KW; w++) g.line("u64 kw" + std::to_string(w) + " = 0;");
        auto put = [&](const std::string& e, int byteOff, int nbytes) {
            // e is a u64 expression already masked to nbytes
            int w = byteOff / 8, sh = (byteOff % 8) * 8;
            g.line("kw" + std::to_string(w) + " |= (" + e + ")" + (sh ? " << " + std::to_string(sh) : "") + ";");
            if (sh + nbytes * 8 > 64) g.line("kw" + std::to_string(w + 1) + " |= (" + e + ") >> " + std::to_string(64 - sh) + ";");
        };
        for (auto& f : fields) {
            const Val& v = pv[f.projChannel];
            if (f.nullable && v.nullable()) put("(u64)(" + v.n + " ? 1 : 0) << " + std::to_string(f.nullBit % 8), f.nullBit / 8, 1);
            std::string lo, hi;
            switch (f.type.kind) {
                case TypeKind::BOOLEAN: lo = "(u64)(" + v.v + " ? 1 : 0)"; break;
                case TypeKind::INTEGER:
                case TypeKind::DATE: lo = "(u64)(u32)" + v.v; break;
                case TypeKind::BIGINT: lo = "(u64)" + v.v; break;
                case TypeKind::DOUBLE: lo = "f64_group_bits(" + v.v + ")"; break;
                case TypeKind::DECIMAL:
                    if (f.type.isShortDecimal()) lo = "(u64)" + v.v;
                    else { lo = "(u64)(u128)" + v.v; hi = "(u64)((u128)" + v.v + " >> 64)"; }
                    break;
                case TypeKind::VARCHAR: lo = v.v; hi = v.hi; break;
                default: unsupported("group key type");
            }
            auto nz = [&](const std::string& e) { return v.nullable() ? "(" + v.n + " ? 0ull : " + e + ")" : e; };
            if (f.nbytes <= 8) {
                put(nz(lo), f.byteOffset, f.nbytes);
            } else {
                put(nz(lo), f.byteOffset, 8);
                put(nz("(" + hi + " & " + maskOf(f.nbytes - 8) + ")"), f.byteOffset + 8, f.nbytes - 8);
            }
        }
        for (int w = 0; w < res.KW; w++) g.line("key[r][" + std::to_string(w) + "] = kw" + std::to_string(w) + ";");
    }

    // ---- cells
    std::vector<Cell> cells;
    std::map<std::string, int> cellIndex;
    int nw = 0;
    auto getCell = [&](const std::string& kind, const std::string& ident, int nwords, const std::vector<std::string>& contrib) -> int {
        std::string k = kind + "|" + ident;
        auto it = cellIndex.find(k);
        if (it != cellIndex.end()) return it->second;
        Cell c;
        c.kind = kind;
        c.word = nw;
        c.nwords = nwords;
        c.contrib = contrib;
        nw += nwords;
        cells.push_back(c);
        cellIndex[k] = (int)cells.size() - 1;
        return (int)cells.size() - 1;
    };
    auto sel1 = [](const std::string& cond, const std::string& e, const std::string& zero) { return cond == "true" ? e : "(" + cond + " ? " + e + " : " + zero + ")"; };

    // min / max over long DECIMAL (AbstractMinMaxAggregationFunction.java:274-306 on Int128 values): ONE 128-bit cell = two words
    // [low, high] of the order-preserving image (two's complement value with the sign bit flipped; complemented for min), updated
    // with a 16-byte compare-and-swap loop. The pair must be 16-byte aligned inside the 32-byte aligned slot (and in the global
    // aggregation's state array): a zero word is put in front when the pair would start at an odd word.
    const bool taglessByKey = res.mode == KernelMode::AGG_HASH && res.KW >= 1 && res.KW <= 2 && keyLen < res.KW * 8 && !getenv("PGPU_TUNE_TAGGED");
    const int cellBaseWord = res.mode == KernelMode::AGG_HASH ? (taglessByKey ? 0 : 1) + res.KW : 0;
    // img: an unsigned 128-bit expression whose order is the values' order
    auto decImage = [](const std::string& v128) { return "((u128)(" + v128 + ") ^ ((u128)1 << 127))"; };
    auto minMax128Cell = [&](bool isMax, const std::string& ident, const std::string& cond, const std::string& img) -> int {
        if ((cellBaseWord + nw) & 1) getCell("pad", "pad" + std::to_string(nw), 1, {"0ull"});
        const std::string enc = isMax ? img : "(~" + img + ")";
        return getCell(isMax ? "max128" : "min128", ident, 2, {sel1(cond, "(u64)" + enc, "0ull"), sel1(cond, "(u64)(" + enc + " >> 64)", "0ull")});
    };
    struct AggCells {
        int cnt = -1, sum = -1, ext = -1;
    };
    std::vector<AggCells> aggCells;
    for (auto& a : plan.aggs) {
        AggCells ac;
        if (!finalStep) {
            std::string cond = "true";
            if (a.mask >= 0) {
                const Val& m = pv[a.mask];
                cond = m.nullable() ? "(" + m.v + " && !" + m.n + ")" : m.v;
            }
            Val x;
            if (a.hasArg) {
                x = pv[a.arg];
                if (x.nullable()) cond = cond == "true" ? "!" + x.n : "(" + cond + " && !" + x.n + ")";
            }
            const int argId = (a.hasArg && mayBeNull(*plan.projections[a.arg])) ? a.arg : -1;
            const std::string cid = "arg=" + std::to_string(argId) + "|mask=" + std::to_string(a.mask);
            const std::string vid = "arg=" + std::to_string(a.arg) + "|mask=" + std::to_string(a.mask);
            ac.cnt = getCell("cnt", cid, 1, {sel1(cond, "1ull", "0ull")});
            const Type& t = a.inType;
            switch (a.fn) {
                case AggFn::COUNT: break;
                case AggFn::SUM:
                case AggFn::AVG:
                    if (t.kind == TypeKind::DOUBLE || (a.fn == AggFn::AVG && t.kind == TypeKind::BIGINT)) {
                        std::string xv = t.kind == TypeKind::DOUBLE ? x.v : "(double)" + x.v;
                        ac.sum = getCell("sumd", vid + (t.kind == TypeKind::DOUBLE ? "" : "|cast"), 1, {"d2u(" + sel1(cond, xv, "0.0") + ")"});
                    } else if (t.kind == TypeKind::BIGINT) {
                        ac.sum = getCell("sumi", vid, 2,
                                         {sel1(cond, "(u64)(u32)" + x.v, "0ull"), sel1(cond, "(u64)(" + x.v + " >> 32)", "0ull")});
                    } else if (t.kind == TypeKind::DECIMAL) {
                        std::string w = "(i128)" + x.v;
                        ac.sum = getCell("sumdec", vid, 4,
                                         {sel1(cond, "(u64)(u32)(u128)(" + w + ")", "0ull"), sel1(cond, "(u64)(u32)((u128)(" + w + ") >> 32)", "0ull"),
                                          sel1(cond, "(u64)(u32)((u128)(" + w + ") >> 64)", "0ull"), sel1(cond, "(u64)(i64)((" + w + ") >> 96)", "0ull")});
                    }
                    break;
                case AggFn::MIN:
                case AggFn::MAX: {
                    bool isMax = a.fn == AggFn::MAX;
                    std::string k;  // signed order key
                    if (t.kind == TypeKind::DOUBLE) {
                        // MinMaxCompare.maxDouble: (value > state) || isNaN(state) => NaN is the smallest for max;
                        // min uses Double.compare (NaN largest)
                        k = isMax ? "((" + x.v + " != " + x.v + ") ? (i64)0x8000000000000000ull : f64_order_key(" + x.v + "))" : "f64_order_key(" + x.v + ")";
                    } else if (t.kind == TypeKind::BOOLEAN) {
                        k = "(i64)(" + x.v + " ? 1 : 0)";
                    } else if (t.isLongDecimal()) {
                        ac.ext = minMax128Cell(isMax, vid, cond, decImage("(i128)" + x.v));
                        break;
                    } else if (t.kind == TypeKind::VARCHAR) {
                        ac.ext = minMax128Cell(isMax, vid, cond, "varchar_order_image(" + x.v + ", " + x.hi + ")");
                        break;
                    } else {
                        k = "(i64)" + x.v;
                    }
                    std::string enc = isMax ? "((u64)" + k + " ^ 0x8000000000000000ull)" : "(~((u64)" + k + " ^ 0x8000000000000000ull))";
                    ac.ext = getCell(isMax ? "max" : "min", vid, 1, {sel1(cond, enc, "0ull")});
                    break;
                }
            }
        } else {
            // FINAL: contributions come from the intermediate channels (AccumulatorCompiler.java:674-767 addIntermediate -> combine)
            const Type& t = a.inType;
            auto nn = [&](const Val& v) { return v.nullable() ? "!" + v.n : std::string("true"); };
            switch (a.fn) {
                case AggFn::COUNT: {
                    const Val& c0 = pv[a.arg];
                    ac.cnt = getCell("cntw", "F" + std::to_string(a.arg), 1, {sel1(nn(c0), "(u64)" + c0.v, "0ull")});
                    break;
                }
                case AggFn::SUM:
                case AggFn::AVG: {
                    if (t.kind == TypeKind::DECIMAL) {
                        // VARBINARY state: [count:8 (avg only)] overflow:8 value:16 (sign-magnitude)
                        const Val& s0 = pv[a.arg];
                        std::string p = s0.v;
                        int o = a.fn == AggFn::AVG ? 1 : 0;
                        std::string cond = nn(s0);
                        std::string q = "((const u64*)" + p + ")";
                        std::string val = "dec_from_raw(raw128{" + q + "[" + std::to_string(o + 1) + "], " + q + "[" + std::to_string(o + 2) + "]})";
                        std::string ov = "(i64)" + q + "[" + std::to_string(o) + "]";
                        if (a.fn == AggFn::AVG) ac.cnt = getCell("cntw", "F" + std::to_string(a.arg), 1, {sel1(cond, q + "[0]", "0ull")});
                        else ac.cnt = getCell("cnt", "F" + std::to_string(a.arg), 1, {sel1(cond, "1ull", "0ull")});
                        ac.sum = getCell("sumdec", "F" + std::to_string(a.arg), 4,
                                         {sel1(cond, "(u64)(u32)(u128)(" + val + ")", "0ull"), sel1(cond, "(u64)(u32)((u128)(" + val + ") >> 32)", "0ull"),
                                          sel1(cond, "(u64)(u32)((u128)(" + val + ") >> 64)", "0ull"),
                                          sel1(cond, "(u64)((i64)((" + val + ") >> 96) + (" + ov + " << 31))", "0ull")});
                    } else if (a.fn == AggFn::AVG) {
                        const Val &s0 = pv[a.arg], &c1 = pv[a.arg + 1];
                        ac.sum = getCell("sumd", "F" + std::to_string(a.arg), 1, {"d2u(" + sel1(nn(s0), s0.v, "0.0") + ")"});
                        ac.cnt = getCell("cntw", "F" + std::to_string(a.arg + 1), 1, {sel1(nn(c1), "(u64)" + c1.v, "0ull")});
                    } else {
                        const Val &c0 = pv[a.arg], &s1 = pv[a.arg + 1];
                        ac.cnt = getCell("cntw", "F" + std::to_string(a.arg), 1, {sel1(nn(c0), "(u64)" + c0.v, "0ull")});
                        if (t.kind == TypeKind::DOUBLE) ac.sum = getCell("sumd", "F" + std::to_string(a.arg + 1), 1, {"d2u(" + sel1(nn(s1), s1.v, "0.0") + ")"});
                        else ac.sum = getCell("sumi", "F" + std::to_string(a.arg + 1), 2, {sel1(nn(s1), "(u64)(u32)" + s1.v, "0ull"), sel1(nn(s1), "(u64)(" + s1.v + " >> 32)", "0ull")});
                    }
                    break;
                }
                case AggFn::MIN:
                case AggFn::MAX: {
                    bool isMax = a.fn == AggFn::MAX;
                    const Val& x = pv[a.arg];
                    std::string cond = nn(x), k;
                    if (t.kind == TypeKind::DOUBLE) k = isMax ? "((" + x.v + " != " + x.v + ") ? (i64)0x8000000000000000ull : f64_order_key(" + x.v + "))" : "f64_order_key(" + x.v + ")";
                    else if (t.kind == TypeKind::BOOLEAN) k = "(i64)(" + x.v + " ? 1 : 0)";
                    else if (t.isLongDecimal() || t.kind == TypeKind::VARCHAR) {
                        ac.cnt = getCell("cnt", "F" + std::to_string(a.arg), 1, {sel1(cond, "1ull", "0ull")});
                        ac.ext = minMax128Cell(isMax, "F" + std::to_string(a.arg), cond, t.kind == TypeKind::VARCHAR ? "varchar_order_image(" + x.v + ", " + x.hi + ")" : decImage("(i128)" + x.v));
                        break;
                    } else k = "(i64)" + x.v;
                    std::string enc = isMax ? "((u64)" + k + " ^ 0x8000000000000000ull)" : "(~((u64)" + k + " ^ 0x8000000000000000ull))";
                    ac.cnt = getCell("cnt", "F" + std::to_string(a.arg), 1, {sel1(cond, "1ull", "0ull")});
                    ac.ext = getCell(isMax ? "max" : "min", "F" + std::to_string(a.arg), 1, {sel1(cond, enc, "0ull")});
                    break;
                }
            }
        }
        aggCells.push_back(ac);
    }
    res.NW = nw;
    for (auto& c : cells)
        for (int j = 0; j < c.nwords; j++)
            res.cellKinds.push_back(c.kind == "sumd" ? 1 : (c.kind == "max" || c.kind == "min") ? 2 : c.kind == "cnt" ? 3 : (c.kind == "max128" || c.kind == "min128") ? 4 + j : 0);
    // hot words: cells with the same kind and the same per-row contribution expression share one word in the scan loop
    std::vector<int> hotOf((size_t)nw, 0), hotKinds;
    {
        std::map<std::string, int> hotIndex;
        for (auto& c : cells) {
            for (int j = 0; j < c.nwords; j++) {
                int w = c.word + j;
                std::string k = std::to_string(res.cellKinds[w]) + "|" + c.contrib[j];
                auto it = hotIndex.find(k);
                if (it == hotIndex.end()) {
                    int h = (int)hotKinds.size();
                    hotIndex[k] = h;
                    hotKinds.push_back(res.cellKinds[w]);
                    g.line("cw[r][" + std::to_string(h) + "] = " + c.contrib[j] + ";");
                    hotOf[w] = h;
                } else {
                    hotOf[w] = it->second;
                }
            }
        }
    }
    std::string rowBody = g.out.str();
    const int NH = (int)hotKinds.size(), NHA = std::max(NH, 1);
    const int NW = res.NW, KW = res.KW;
    const int NWA = std::max(NW, 1), KWA = std::max(KW, 1);
    // tagless slots: keys of 1-2 words whose last word has a spare top byte carry the "occupied" mark themselves
    const bool tagless = res.mode == KernelMode::AGG_HASH && KW >= 1 && KW <= 2 && keyLen < KW * 8 && !getenv("PGPU_TUNE_TAGGED");
    res.SLOTW = (((tagless ? 0 : 1) + KW + NW) + 3) / 4 * 4;

    // directory sizing (shared memory): 8 warps x DCAP x NW words + keys + tags
    res.useDir = res.mode == KernelMode::AGG_HASH && plan.expectedGroups <= (1 << 20) && NW > 0;
    res.dirCap = 0;
    if (res.useDir) {
        int cap = 128;
        const int nwarps = stagedPossible ? kStagedWarps : 8;  // the staged variant keeps one warp-private table per consumer warp
        while (cap > 16 && (nwarps * cap * NHA * 8 + cap * KW * 8 + cap * 12 + 16) > 40 * 1024) cap /= 2;
        res.dirCap = cap;
        res.smemBytes = 8 * cap * NHA * 8 + cap * KW * 8 + cap * 12 + 16;
    }

    // high cardinality (no directory): run compaction through warp-private shared memory when the staged rows of 8 warps fit
    // in the 48 KB every kernel may use; placement keeps 2^locB neighbouring keys of an integer first key field in
    // consecutive slots
    const int runsBytes = 8 * ((KW + NHA) * 128 * 8 + 256);
    const bool highCard = res.mode == KernelMode::AGG_HASH && !res.useDir && tagless;
    const bool runs = highCard && runsBytes <= 48 * 1024 && envInt("PGPU_TUNE_RUNS", plan.step == Step::SINGLE ? kDefaultRunsSingleStep : plan.step == Step::PARTIAL ? kDefaultRunsPartialStep : kDefaultRunsOtherSteps) != 0;
    if (runs) res.smemBytes = runsBytes;
    // bucketed table: one integer key (so that "the 3 low bits" means neighbouring keys), two spare bytes in the last key word
    // (presence byte + occupied mark), run compaction available (the bucket path is built on it, whatever the step)
    // (the staging bound uses NW, not this signature's NH: whether a plan is bucketed must not depend on which pages carry nulls)
    const int runsBytesAnySig = 8 * ((KW + NWA) * 128 * 8 + 256);
    bool bucketsOk = highCard && fields.size() == 1 && runsBytesAnySig <= 64 * 1024 && NW > 0 && keyLen <= KW * 8 - 2 && fields[0].byteOffset < 7;
    for (int k : res.cellKinds)
        if (k >= 4) bucketsOk = false;  // 128-bit cells need their alignment inside the slot: one key per slot only
    if (bucketsOk) {
        const TypeKind k0 = fields[0].type.kind;
        bucketsOk = k0 == TypeKind::BIGINT || k0 == TypeKind::INTEGER || k0 == TypeKind::DATE || (k0 == TypeKind::DECIMAL && fields[0].type.isShortDecimal());
    }
    res.bucketsPossible = bucketsOk;
    res.buckets = bucketsOk && opt.buckets;
    if (bucketsOk) {
        const Expr& ke = *plan.projections[fields[0].projChannel];
        if (ke.kind == ExprKind::INPUT) res.bucketKeyChannel = ke.channel;
    }
    const int bslotw = (KW + 8 * NW + 3) / 4 * 4;
    if (res.buckets) {
        res.SLOTW = bslotw;
        res.smemBytes = runsBytes;
    }
    int locB = 0, locSh = 0;
    if (highCard && !res.buckets && !fields.empty() && fields[0].byteOffset < 7) {
        const TypeKind k0 = fields[0].type.kind;
        if (k0 == TypeKind::BIGINT || k0 == TypeKind::INTEGER || k0 == TypeKind::DATE || (k0 == TypeKind::DECIMAL && fields[0].type.isShortDecimal())) {
            locB = std::min(std::max(envInt("PGPU_TUNE_LOCB", kDefaultLocBits), 0), 6);
            locSh = 8 * fields[0].byteOffset;
        }
    }
    res.locBits = locB;

    // staged variant: low-cardinality hash aggregation whose stage ring (>= 2 stages) fits beside the directory
    {
        const int kBudget = 220 * 1024;
        int cap = std::max(res.dirCap, 1);
        int dirBytes = res.useDir ? kStagedWarps * cap * NHA * 8 + cap * KW * 8 + cap * 12 + 16 : 0;
        res.stagedTile = kTile;
        res.stagedThreads = (kStagedWarps + 1) * 32;
        const int dcBytes = g.dcSlot.empty() ? 0 : (Gen::kDcHeader + (int)g.dcSlot.size() * Gen::kDcSlotBytes + 127) / 128 * 128;
        res.stagedFixedSmem = (dirBytes + 127) / 128 * 128 + 128 + dcBytes;
        res.dcacheCap = Gen::kDcCap;
        for (auto& kv : g.dcSlot) res.dcacheChannels.push_back(kv.first);
        res.stageBytes = stageBytes;
        res.staged = stagedPossible && res.useDir && !getenv("PGPU_TUNE_NOSTAGE") && stageBytes > 0 && res.stagedFixedSmem + 2 * stageBytes <= kBudget;
    }

    // ---- outputs
    std::vector<Type> outTypes = plan.outputTypes();
    res.NO = (int)outTypes.size();
    for (auto& t : outTypes) {
        OutColDesc d;
        d.type = t;
        if (t.kind == TypeKind::VARCHAR) {
            d.stride = 16;
            d.varWidth = true;
            d.lenPrefixed = true;
        } else if (t.kind == TypeKind::VARBINARY) {
            d.varWidth = true;
            d.stride = 0;  // fixed below per aggregate
        } else {
            d.stride = t.width();
        }
        res.out.push_back(d);
    }

    src << "struct P {\n";
    src << "    static constexpr int NC = " << res.NC << ", NO = " << res.NO << ", KW = " << KW << ", NW = " << NW << ", SLOTW = " << res.SLOTW << ";\n";
    src << "    static constexpr int KWA = " << KWA << ", NWA = " << NWA << ", NH = " << NH << ", NHA = " << NHA << ";\n";
    src << "    static constexpr bool TAGLESS = " << (tagless ? "true" : "false") << ";\n";
    src << "    static constexpr bool LANE_MERGE = " << (getenv("PGPU_TUNE_NOLANEMERGE") ? "false" : "true") << ";\n";
    src << "    static constexpr bool RUNS = " << (runs ? "true" : "false") << ";  // run compaction (HashCtx::consume_runs)\n";
    src << "    static constexpr bool BUCKETS = " << (res.buckets ? "true" : "false") << ";  // bucketed table: SLOTW words per bucket of 8 keys\n";
    src << "    static constexpr int BSLOTW = " << bslotw << ", BSH = " << (fields.empty() ? 0 : 8 * fields[0].byteOffset) << ", BPRES = 48;\n";
    src << "    static constexpr u64 BPMASK = 0xffull << BPRES;\n";
    {
        // range-aware exchange split (body_materialize_range): one integer grouping key in tagless one-key-per-slot tables
        bool rangeOk = res.mode == KernelMode::AGG_HASH && tagless && !res.buckets && fields.size() == 1 && NW > 0 && fields[0].byteOffset < 7;
        if (rangeOk) {
            const TypeKind k0 = fields[0].type.kind;
            rangeOk = k0 == TypeKind::BIGINT || k0 == TypeKind::INTEGER || k0 == TypeKind::DATE || (k0 == TypeKind::DECIMAL && fields[0].type.isShortDecimal());
        }
        res.rangeOk = rangeOk;
        unsigned long long nullMask = 0;
        if (rangeOk && fields[0].nullable) nullMask = 1ull << fields[0].nullBit;
        res.rangeNullable = nullMask != 0;
        src << "    static constexpr bool RANGE_OK = " << (rangeOk ? "true" : "false") << ";\n";
        src << "    static constexpr u64 BNULLMASK = " << nullMask << "ull;  // key word 0: the grouping key is NULL\n";
    }
    src << "    static constexpr bool BCASFIRST = " << (envInt("PGPU_TUNE_BCAS", 1) ? "true" : "false") << ";  // bucket leaders claim with one compare-and-swap, no load first\n";
    src << "    static constexpr int BKBITS = " << (fields.empty() ? 64 : std::min(64, 8 * fields[0].nbytes)) << ";  // width of the (first) key field\n";
    src << "    static constexpr int KOFF = " << (tagless ? 0 : 1) << ";  // first key word of a slot\n";
    src << "    static constexpr bool USE_DIR = " << (res.useDir ? "true" : "false") << ";\n";
    src << "    static constexpr int DCAP = " << std::max(res.dirCap, 1) << ", DLIMIT = " << std::max(res.dirCap * 3 / 4, 1) << ", NWARP = 8, MAX_WARP_GROUPS = 8;\n";
    // register-resident groups: as many as fit in ~28 64-bit accumulator registers per thread, at most 4
    int rg = res.useDir ? std::min(4, 28 / NHA) : 0;
    if (const char* e = getenv("PGPU_TUNE_RG")) rg = res.useDir ? std::min(atoi(e), 8) : 0;  // tuning knob (bench sweeps)
    src << "    static constexpr int RG = " << rg << ";\n";
    src << "    static constexpr int EAGER_LANES = " << (getenv("PGPU_TUNE_EAGER") ? atoi(getenv("PGPU_TUNE_EAGER")) : 16) << ";\n";
    auto kindName = [&](int k) { return k == 0 ? "CK_ADD_U64" : k == 1 ? "CK_ADD_F64" : k == 2 ? "CK_MAX_U64" : "CK_CNT"; };
    src << "    PGPU_DEV static void identity(u64 (&a)[NHA]) {\n";
    for (int w = 0; w < NHA; w++) src << "        a[" << w << "] = 0ull;\n";
    src << "    }\n";
    src << "    PGPU_DEV static void identity_ptr(u64* a) {\n";
    for (int w = 0; w < NH; w++) src << "        a[" << w << "] = 0ull;\n";
    src << "    }\n";
    // kinds 4 / 5 = the low / high word of a 128-bit maximum (min / max over long DECIMAL): handled as a pair at the low word
    auto emitCombine = [&](const char* name, const char* sigA, const char* sigB) {
        src << "    PGPU_DEV static void " << name << "(" << sigA << ", " << sigB << ") {\n";
        for (int w = 0; w < NH; w++) {
            if (hotKinds[w] == 4) src << "        max128_combine(a[" << w << "], a[" << w + 1 << "], b[" << w << "], b[" << w + 1 << "]);\n";
            else if (hotKinds[w] != 5) src << "        a[" << w << "] = cell_combine<" << kindName(hotKinds[w]) << ">(a[" << w << "], b[" << w << "]);\n";
        }
        src << "    }\n";
    };
    emitCombine("combine", "u64 (&a)[NHA]", "const u64 (&b)[NHA]");
    emitCombine("combine_ptr", "u64* a", "const u64 (&b)[NHA]");
    emitCombine("combine_ptr_to", "u64 (&a)[NHA]", "const u64* b");
    src << "    PGPU_DEV static void warp_reduce_cells(u64 (&a)[NHA]) {\n";
    for (int w = 0; w < NH; w++) {
        if (hotKinds[w] == 4) src << "        max128_warp_reduce(a[" << w << "], a[" << w + 1 << "]);\n";
        else if (hotKinds[w] != 5) src << "        a[" << w << "] = warp_reduce<" << kindName(hotKinds[w]) << ">(a[" << w << "]);\n";
    }
    src << "    }\n";
    // hot -> cells expansion: atomically into a table slot, or plainly into the global-aggregation state
    src << "    PGPU_DEV static void atomic_cells(u64* cells, const u64 (&hot)[NHA]) {\n";
    for (int w = 0; w < NW; w++) {
        if (res.cellKinds[w] == 4) src << "        max128_atomic(cells + " << w << ", hot[" << hotOf[w] << "], hot[" << hotOf[w + 1] << "]);\n";
        else if (res.cellKinds[w] != 5) src << "        cell_atomic<" << kindName(res.cellKinds[w]) << ">(cells + " << w << ", hot[" << hotOf[w] << "]);\n";
    }
    src << "    }\n";
    src << "    PGPU_DEV static void merge_cells(u64* cells, const u64 (&hot)[NHA]) {\n";
    for (int w = 0; w < NW; w++) {
        if (res.cellKinds[w] == 4) src << "        max128_combine(cells[" << w << "], cells[" << w + 1 << "], hot[" << hotOf[w] << "], hot[" << hotOf[w + 1] << "]);\n";
        else if (res.cellKinds[w] != 5) src << "        cells[" << w << "] = cell_combine<" << kindName(res.cellKinds[w]) << ">(cells[" << w << "], hot[" << hotOf[w] << "]);\n";
    }
    src << "    }\n";

    // cells (+) cells: merging two partial states of one group (streaming runs: pieces of a run, rows into the table)
    src << "    PGPU_DEV static void combine_cells(u64* a, const u64* b) {\n";
    for (int w = 0; w < NW; w++) {
        if (res.cellKinds[w] == 4) src << "        max128_combine(a[" << w << "], a[" << w + 1 << "], b[" << w << "], b[" << w + 1 << "]);\n";
        else if (res.cellKinds[w] != 5) src << "        a[" << w << "] = cell_combine<" << kindName(res.cellKinds[w]) << ">(a[" << w << "], b[" << w << "]);\n";
    }
    src << "    }\n";
    src << "    PGPU_DEV static void atomic_cells_from(u64* cells, const u64* b) {\n";
    for (int w = 0; w < NW; w++) {
        if (res.cellKinds[w] == 4) src << "        max128_atomic(cells + " << w << ", b[" << w << "], b[" << w + 1 << "]);\n";
        else if (res.cellKinds[w] != 5) src << "        cell_atomic<" << kindName(res.cellKinds[w]) << ">(cells + " << w << ", b[" << w << "]);\n";
    }
    src << "    }\n";

    // ---- eval
    // eval: filter channels first, the remaining channels only when some row of the thread passed (selective filters
    // never read them: Q6 skips most of extendedprice). eval_eager: every load issued up front, one memory round trip
    // per tile (the kernels switch to it while most lanes pass the filter: Q1).
    for (int eager = 0; eager < 2; eager++) {
        src << "    PGPU_DEV static void " << (eager ? "eval_eager" : "eval") << "(const PageArgs<NC>& a, i64 base, i64 nrows, bool (&sel)[4], u64 (&key)[4][KWA], u64 (&cw)[4][NHA], u32& err) {\n";
        src << (eager ? evalFilterEager.str() : evalFilter.str());
        src << "        #pragma unroll\n";
        src << "        for (int r = 0; r < 4; r++) {\n";
        src << "            if (sel[r]) {\n";
        src << rowPrologue << rowBody;
        src << "            }\n";
        src << "        }\n";
        src << "    }\n";
    }
    (void)projRaises;
    if (res.staged) {
        src << "    static constexpr int SR = " << g.SR << ", CWARPS = " << kStagedWarps << ", SUB = " << kSub << ", TILE = " << kTile << ", STAGE_BYTES = " << stageBytes << ", TAIL_GUARD = " << tailGuard << ";\n";
        {
            const int dcBytes = g.dcSlot.empty() ? 0 : (Gen::kDcHeader + (int)g.dcSlot.size() * Gen::kDcSlotBytes + 127) / 128 * 128;
            src << "    static constexpr int DCACHE_BYTES = " << dcBytes << ";\n";
            // executed by the whole block before the ring starts; the caller synchronises afterwards
            src << "    PGPU_DEV static void dcache_fill(const PageArgs<NC>& a, unsigned char* dc) {\n";
            // (the host only launches the staged kernel when every cached dictionary has at most kDcCap entries)
            for (auto& kv : g.dcSlot) {
                const int c = kv.first, k = kv.second;
                const Type& t = plan.input[c];
                const std::string C = std::to_string(c);
                const int voff = Gen::kDcHeader + k * Gen::kDcSlotBytes, noff = voff + Gen::kDcCap * 8;
                src << "        for (int i = threadIdx.x; i < a.c[" << C << "].dcount && i < " << Gen::kDcCap << "; i += blockDim.x) {\n";
                if (t.kind == TypeKind::VARCHAR) {
                    src << "            u64 x, y; u32 e = 0; const i32 q0 = a.c[" << C << "].doffsets[i], q1 = a.c[" << C << "].doffsets[i + 1];\n";
                    src << "            pack_varchar<" << t.precision << ">((const u8*)a.c[" << C << "].dvalues, q0, q1 - q0, x, y, e);\n";
                    src << "            ((u64*)(dc + " << voff << "))[i] = e ? 0xffull : x;\n";
                } else {
                    const std::string raw = "((const " + stype(t) + "*)a.c[" + C + "].dvalues)[i]";
                    src << "            ((u64*)(dc + " << voff << "))[i] = " << (t.kind == TypeKind::DOUBLE ? "d2u(" + raw + ")" : "(u64)(i64)" + raw) << ";\n";
                }
                if (sig[c].nullable) src << "            (dc + " << noff << ")[i] = a.c[" << C << "].dnulls[i];\n";
                src << "        }\n";
            }
            src << "    }\n";
        }
        src << "    PGPU_DEV static void eval_staged(const PageArgs<NC>& a, const unsigned char* st, const unsigned char* dc, int trow, bool (&sel)[SR], u64 (&key)[SR][KWA], u64 (&cw)[SR][NHA], u32& err) {\n";
        src << evalFilterStaged.str();
        src << "        #pragma unroll\n";
        src << "        for (int r = 0; r < SR; r++) {\n";
        src << "            if (sel[r]) {\n";
        src << rowPrologueStaged << rowBody;
        src << "            }\n";
        src << "        }\n";
        src << "    }\n";
        // producer warp: arm the stage's full barrier with the byte count, then hand the copies to the TMA engine
        src << "    struct StagePre {\n        int unused;\n" << preFields.str() << "    };\n";
        src << "    struct StageLim {\n        int unused;\n" << limFields.str() << "    };\n";
        src << "    PGPU_DEV static void stage_prefetch(const PageArgs<NC>& a, i64 row0, StagePre& p) {\n        p.unused = 0;\n" << preLoads.str() << "    }\n";
        src << "    PGPU_DEV static void stage_limits(const PageArgs<NC>& a, StageLim& l) {\n        l.unused = 0;\n" << limLoads.str() << "    }\n";
        src << "    PGPU_DEV static void stage_issue(const PageArgs<NC>& a, i64 row0, unsigned char* st, u64* bar, int lane, u32& err, u64* empty_bar, bool wait_empty, u32 empty_parity, const StagePre& p, const StageLim& l) {\n";
        src << "        u32 tx = " << fixedTx << "u;\n";
        src << issueLoad.str();
        src << "        if (wait_empty) mbar_wait(empty_bar, empty_parity);\n";
        src << issueVar.str();
        src << "        __syncwarp();\n";
        src << "        if (lane == 0) {\n";
        src << "            mbar_arrive_expect_tx(bar, tx);\n";
        src << issueCopies.str();
        src << "        }\n";
        src << "    }\n";
    }

    // ---- emit (finalisation of one group)
    {
        std::ostringstream em;
        auto L = [&](const std::string& s) { em << "        " << s << "\n"; };
        int oc = 0;
        // keys
        for (auto& f : fields) {
            std::string O = std::to_string(oc);
            auto get = [&](int byteOff, int nbytes) -> std::string {
                int w = byteOff / 8, sh = (byteOff % 8) * 8;
                std::string e = "(key[" + std::to_string(w) + "]" + (sh ? " >> " + std::to_string(sh) : "") + ")";
                if (sh + nbytes * 8 > 64) e = "(" + e + " | (key[" + std::to_string(w + 1) + "] << " + std::to_string(64 - sh) + "))";
                if (nbytes < 8) e = "(" + e + " & " + maskOf(nbytes) + ")";
                return e;
            };
            if (f.nullable) L("out.c[" + O + "].nulls[pos] = (u8)(" + get(f.nullBit / 8, 1) + " >> " + std::to_string(f.nullBit % 8) + ") & 1;");
            else L("out.c[" + O + "].nulls[pos] = 0;");
            switch (f.type.kind) {
                case TypeKind::BOOLEAN: L("((u8*)out.c[" + O + "].values)[pos] = (u8)" + get(f.byteOffset, 1) + ";"); break;
                case TypeKind::INTEGER:
                case TypeKind::DATE: L("((i32*)out.c[" + O + "].values)[pos] = (i32)(u32)" + get(f.byteOffset, 4) + ";"); break;
                case TypeKind::BIGINT:
                case TypeKind::DOUBLE: L("((u64*)out.c[" + O + "].values)[pos] = " + get(f.byteOffset, 8) + ";"); break;
                case TypeKind::DECIMAL:
                    if (f.type.isShortDecimal()) L("((u64*)out.c[" + O + "].values)[pos] = " + get(f.byteOffset, 8) + ";");
                    else L("((raw128*)out.c[" + O + "].values)[pos] = dec_to_raw((i128)(((u128)" + get(f.byteOffset + 8, 8) + " << 64) | (u128)" + get(f.byteOffset, 8) + "));");
                    break;
                case TypeKind::VARCHAR: {
                    std::string lo = get(f.byteOffset, std::min(8, f.nbytes));
                    std::string hi = f.nbytes > 8 ? get(f.byteOffset + 8, f.nbytes - 8) : "0ull";
                    L("{ raw128 q; q.lo = " + lo + "; q.hi = " + hi + "; ((raw128*)out.c[" + O + "].values)[pos] = q; }");
                    break;
                }
                default: break;
            }
            oc++;
        }
        // aggregates
        auto cw = [&](int cell, int j) { return "cells[" + std::to_string(cells[cell].word + j) + "]"; };
        auto sum128 = [&](int cell) {  // 4 limbs of 32 bits (top limb signed 64) -> i128 + fits flag in `fit`
            std::string c0 = cw(cell, 0), c1 = cw(cell, 1), c2 = cw(cell, 2), c3 = cw(cell, 3);
            L("i128 s128; bool fit;");
            L("{ u128 t0 = (u128)" + c0 + " + ((u128)" + c1 + " << 32) + ((u128)" + c2 + " << 64); i128 hi = (i128)(i64)" + c3 + " + (i128)(u64)(t0 >> 96);");
            L("  fit = (hi >= -((i128)1 << 31)) && (hi < ((i128)1 << 31)); s128 = (i128)(((u128)hi << 96) | (t0 & (((u128)1 << 96) - 1))); }");
        };
        for (size_t ai = 0; ai < plan.aggs.size(); ai++) {
            const AggSpec& a = plan.aggs[ai];
            const AggCells& ac = aggCells[ai];
            const Type& t = a.inType;
            std::string cnt = ac.cnt >= 0 ? cw(ac.cnt, 0) : "0ull";
            auto O = [&](int k) { return std::to_string(oc + k); };
            auto setNull = [&](int k, const std::string& cond) { L("out.c[" + O(k) + "].nulls[pos] = (" + cond + ") ? 1 : 0;"); };
            L("{");
            auto extDecode = [&](bool isMax) -> std::string {
                std::string raw = cw(ac.ext, 0);
                return isMax ? "(i64)(" + raw + " ^ 0x8000000000000000ull)" : "(i64)((~" + raw + ") ^ 0x8000000000000000ull)";
            };
            auto storeMinMax = [&](int k, bool isMax) {
                if (t.kind == TypeKind::VARCHAR) {  // the image back to the 16-byte cell [length | bytes]
                    const std::string img = "(((u128)" + cw(ac.ext, 1) + " << 64) | (u128)" + cw(ac.ext, 0) + ")";
                    L("((raw128*)out.c[" + O(k) + "].values)[pos] = varchar_from_order_image(" + std::string(isMax ? "" : "~") + img + ");");
                    return;
                }
                if (t.isLongDecimal()) {  // the image back to the value, then the block's sign-magnitude words
                    const std::string img = "(((u128)" + cw(ac.ext, 1) + " << 64) | (u128)" + cw(ac.ext, 0) + ")";
                    L("{ raw128 q = dec_to_raw((i128)(" + std::string(isMax ? "" : "~") + img + " ^ ((u128)1 << 127))); u64* o = (u64*)out.c[" + O(k) + "].values + 2 * pos; o[0] = q.lo; o[1] = q.hi; }");
                    return;
                }
                std::string kx = extDecode(isMax);
                if (t.kind == TypeKind::DOUBLE) {
                    if (isMax) L("{ i64 kx = " + kx + "; ((double*)out.c[" + O(k) + "].values)[pos] = (kx == (i64)0x8000000000000000ull) ? u2d(0x7ff8000000000000ull) : f64_from_order_key(kx); }");
                    else L("((double*)out.c[" + O(k) + "].values)[pos] = f64_from_order_key(" + kx + ");");
                } else if (t.kind == TypeKind::BOOLEAN) L("((u8*)out.c[" + O(k) + "].values)[pos] = (u8)" + kx + ";");
                else if (t.isInt32()) L("((i32*)out.c[" + O(k) + "].values)[pos] = (i32)" + kx + ";");
                else L("((i64*)out.c[" + O(k) + "].values)[pos] = " + kx + ";");
            };
            const bool partial = outputIsIntermediate(plan.step);  // PARTIAL / INTERMEDIATE: states go out
            switch (a.fn) {
                case AggFn::COUNT:
                    L("((i64*)out.c[" + O(0) + "].values)[pos] = (i64)" + cnt + ";");
                    setNull(0, "false");
                    oc += 1;
                    break;
                case AggFn::SUM:
                case AggFn::AVG:
                    if (t.kind == TypeKind::DECIMAL) {
                        sum128(ac.sum);
                        if (partial) {
                            // VARBINARY [count] overflow value (LongDecimalWithOverflow[AndLong]StateSerializer); canonical form: overflow = 0 when the sum fits
                            int stride = a.fn == AggFn::AVG ? 32 : 24;
                            res.out[oc].stride = stride;
                            std::string base = "((u64*)((u8*)out.c[" + O(0) + "].values + pos * " + std::to_string(stride) + "))";
                            L("{ i64 ovf = 0; i128 v = s128;");
                            L("  if (!fit) { /* |sum| >= 2^127: split into overflow * 2^127 + remainder */");
                            std::string c3 = cw(ac.sum, 3);
                            L("    u128 t0 = (u128)" + cw(ac.sum, 0) + " + ((u128)" + cw(ac.sum, 1) + " << 32) + ((u128)" + cw(ac.sum, 2) + " << 64); i128 hi = (i128)(i64)" + c3 + " + (i128)(u64)(t0 >> 96);");
                            L("    ovf = (i64)(hi >> 31); i128 rem = hi - ((i128)ovf << 31); v = (i128)(((u128)rem << 96) | (t0 & (((u128)1 << 96) - 1))); }");
                            L("  raw128 q = dec_to_raw(v); u64* o = " + base + ";");
                            if (a.fn == AggFn::AVG) L("  o[0] = " + cnt + "; o[1] = (u64)ovf; o[2] = q.lo; o[3] = q.hi; }");
                            else L("  o[0] = (u64)ovf; o[1] = q.lo; o[2] = q.hi; }");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        } else if (a.fn == AggFn::SUM) {
                            // DecimalSumAggregation.outputLongDecimal :177-189
                            L("if (" + cnt + " != 0ull && (!fit || dec_overflows(s128))) err |= PGPU_ERR_OVERFLOW;");
                            L("((raw128*)out.c[" + O(0) + "].values)[pos] = dec_to_raw(s128);");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        } else {
                            // DecimalAverageAggregation.average :218-227: sum / count, HALF_UP at the type's scale
                            L("i128 avg = 0; if (" + cnt + " != 0ull) { if (!fit) err |= PGPU_ERR_OVERFLOW; else avg = dec_div_round(s128, (i128)" + cnt + "); }");
                            if (a.outType.isShortDecimal()) L("((i64*)out.c[" + O(0) + "].values)[pos] = (i64)avg;");
                            else L("((raw128*)out.c[" + O(0) + "].values)[pos] = dec_to_raw(avg);");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        }
                    } else if (a.fn == AggFn::AVG) {
                        std::string s = "u2d(" + cw(ac.sum, 0) + ")";
                        if (partial) {  // ROW(double sum, bigint count)
                            L("((double*)out.c[" + O(0) + "].values)[pos] = " + s + "; ((i64*)out.c[" + O(1) + "].values)[pos] = (i64)" + cnt + ";");
                            setNull(0, "false");
                            setNull(1, "false");
                            oc += 2;
                        } else {  // AverageAggregations.output :66-79
                            L("((double*)out.c[" + O(0) + "].values)[pos] = " + cnt + " ? " + s + " / (double)(i64)" + cnt + " : 0.0;");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        }
                    } else if (t.kind == TypeKind::DOUBLE) {
                        std::string s = "u2d(" + cw(ac.sum, 0) + ")";
                        if (partial) {  // ROW(bigint count, double sum)
                            L("((i64*)out.c[" + O(0) + "].values)[pos] = (i64)" + cnt + "; ((double*)out.c[" + O(1) + "].values)[pos] = " + s + ";");
                            setNull(0, "false");
                            setNull(1, "false");
                            oc += 2;
                        } else {  // DoubleSumAggregation.output :54-62
                            L("((double*)out.c[" + O(0) + "].values)[pos] = " + s + ";");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        }
                    } else {  // sum(bigint): LongSumAggregation — exact; NUMERIC_VALUE_OUT_OF_RANGE when the total leaves int64
                        L("i128 tot = (i128)(u128)" + cw(ac.sum, 0) + " + ((i128)(i64)" + cw(ac.sum, 1) + " << 32);");
                        L("if (tot > (i128)0x7fffffffffffffffll || tot < -(i128)0x7fffffffffffffffll - 1) err |= PGPU_ERR_OVERFLOW;");
                        if (partial) {
                            L("((i64*)out.c[" + O(0) + "].values)[pos] = (i64)" + cnt + "; ((i64*)out.c[" + O(1) + "].values)[pos] = (i64)tot;");
                            setNull(0, "false");
                            setNull(1, "false");
                            oc += 2;
                        } else {
                            L("((i64*)out.c[" + O(0) + "].values)[pos] = (i64)tot;");
                            setNull(0, cnt + " == 0ull");
                            oc += 1;
                        }
                    }
                    break;
                case AggFn::MIN:
                case AggFn::MAX:
                    storeMinMax(0, a.fn == AggFn::MAX);
                    setNull(0, cnt + " == 0ull");
                    oc += 1;
                    break;
            }
            L("}");
        }
        src << "    PGPU_DEV static void emit(const u64* key, const u64* cells, const OutArgs<NO>& out, i64 pos, u32& err) {\n";
        src << em.str();
        src << "    }\n";
    }
    // ---- key_order (streaming runs): the order the rows of a run buffer are expected in = the order of a table sorted by the
    // grouping fields, first field first, every field as its SQL type orders it (signed integers, doubles by value), NULLs last.
    // < 0: a before b, 0: the same key, > 0: a after b. The words are taken without the slot's occupied mark.
    if (res.mode == KernelMode::AGG_HASH && !fields.empty()) {
        std::ostringstream ko;
        auto L = [&](const std::string& s) { ko << "        " << s << "\n"; };
        auto getOf = [&](const std::string& arr, int byteOff, int nbytes) -> std::string {
            int w = byteOff / 8, sh = (byteOff % 8) * 8;
            std::string e = "(" + arr + "[" + std::to_string(w) + "]" + (sh ? " >> " + std::to_string(sh) : "") + ")";
            if (sh + nbytes * 8 > 64) e = "(" + e + " | (" + arr + "[" + std::to_string(w + 1) + "] << " + std::to_string(64 - sh) + "))";
            if (nbytes < 8) e = "(" + e + " & " + maskOf(nbytes) + ")";
            return e;
        };
        for (auto& f : fields) {
            L("{");
            if (f.nullable) {
                L("    const u32 na = (u32)(" + getOf("a", f.nullBit / 8, 1) + " >> " + std::to_string(f.nullBit % 8) + ") & 1u, nb = (u32)(" + getOf("b", f.nullBit / 8, 1) + " >> " +
                  std::to_string(f.nullBit % 8) + ") & 1u;");
                L("    if (na != nb) return na < nb ? -1 : 1;");
            }
            const int n0 = std::min(8, f.nbytes);
            std::string la = getOf("a", f.byteOffset, n0), lb = getOf("b", f.byteOffset, n0);
            switch (f.type.kind) {
                case TypeKind::INTEGER:
                case TypeKind::DATE:
                    L("    const i32 x = (i32)(u32)" + la + ", y = (i32)(u32)" + lb + ";");
                    break;
                case TypeKind::BIGINT:
                    L("    const i64 x = (i64)" + la + ", y = (i64)" + lb + ";");
                    break;
                case TypeKind::DOUBLE:  // the group bits of a double as a sign-magnitude integer: -x before +x, NaN after everything
                    L("    const u64 xa = " + la + ", xb = " + lb + ";");
                    L("    const u64 x = (xa >> 63) ? ~xa : (xa | 0x8000000000000000ull), y = (xb >> 63) ? ~xb : (xb | 0x8000000000000000ull);");
                    break;
                case TypeKind::DECIMAL:
                    if (f.type.isShortDecimal()) {
                        L("    const i64 x = (i64)" + la + ", y = (i64)" + lb + ";");
                    } else {  // 16 value bytes: the high half decides (signed), then the low half
                        L("    const i64 xh = (i64)" + getOf("a", f.byteOffset + 8, f.nbytes - 8) + ", yh = (i64)" + getOf("b", f.byteOffset + 8, f.nbytes - 8) + ";");
                        L("    if (xh != yh) return xh < yh ? -1 : 1;");
                        L("    const u64 x = " + la + ", y = " + lb + ";");
                    }
                    break;
                default:  // BOOLEAN, packed VARCHAR: any fixed order of the bytes
                    if (f.nbytes > 8) {
                        L("    const u64 xh = " + getOf("a", f.byteOffset + 8, f.nbytes - 8) + ", yh = " + getOf("b", f.byteOffset + 8, f.nbytes - 8) + ";");
                        L("    if (xh != yh) return xh < yh ? -1 : 1;");
                    }
                    L("    const u64 x = " + la + ", y = " + lb + ";");
                    break;
            }
            L("    if (x != y) return x < y ? -1 : 1;");
            L("}");
        }
        L("return 0;");
        src << "    PGPU_DEV static int key_order(const u64* a, const u64* b) {\n" << ko.str() << "    }\n";
        const Expr& ke = *plan.projections[fields[0].projChannel];
        const TypeKind k0 = fields[0].type.kind;
        if (ke.kind == ExprKind::INPUT && (k0 == TypeKind::BIGINT || k0 == TypeKind::INTEGER || k0 == TypeKind::DATE || (k0 == TypeKind::DECIMAL && fields[0].type.isShortDecimal())))
            res.streamKeyChannel = ke.channel;
    }
    // partition sweep (device header, "partition sweep"): the rows of the used channels can be moved one by one when every
    // array is fixed width per position: plain values / 16-byte varchar cells / dictionary ids (+ null flags); RLE blocks have
    // nothing to move
    bool sweepOk = highCard && !res.buckets && res.mode == KernelMode::AGG_HASH && plan.maxPartialGroups == 0 && envInt("PGPU_TUNE_SWEEP", 1) != 0;
    int sweepRowBytes = 0, sweepStageWidth = 1;
    {
        // move_tile: one moved array at a time, the tile's elements are staged in shared memory with coalesced loads (in row order)
        // and written out in slice order — every global byte is read once and written once, whatever the order of the keys
        std::ostringstream mv;
        auto moveArray = [&](const std::string& srcExpr, const std::string& dstExpr, int w) {
            const char* ty = w == 1 ? "u8" : w == 4 ? "u32" : "u64";
            const int words = w == 16 ? 2 : 1;  // 16-byte values as two words: the caller's arrays need not be 16-byte aligned
            mv << "        {\n            " << ty << "* sm = (" << ty << "*)stage;\n";
            mv << "            const " << ty << "* src = (const " << ty << "*)(" << srcExpr << ") + tbase * " << words << ";\n";
            mv << "            " << ty << "* dst = (" << ty << "*)const_cast<void*>((const void*)(" << dstExpr << "));\n";
            mv << "            for (int i = threadIdx.x; i < nt * " << words << "; i += blockDim.x) sm[i] = src[i];\n";
            mv << "            __syncthreads();\n";
            mv << "            for (u32 j = threadIdx.x; j < total; j += blockDim.x) {\n";
            mv << "                const u32 pp = spid[j];\n                const u64 d = gbase[pp] + (j - lstart[pp]);\n                const u32 sr = sidx[j];\n";
            if (words == 2) mv << "                dst[2 * d] = sm[2 * sr];\n                dst[2 * d + 1] = sm[2 * sr + 1];\n";
            else mv << "                dst[d] = sm[sr];\n";
            mv << "            }\n            __syncthreads();\n        }\n";
            sweepRowBytes += w;
            sweepStageWidth = std::max(sweepStageWidth, w);
        };
        for (int c : res.usedChannels) {
            const Type& t = plan.input[c];
            const ColSig& cs = sig[c];
            const std::string C = std::to_string(c);
            if (cs.encoding == PGPU_ENC_RLE) continue;
            if (cs.encoding == PGPU_ENC_DICTIONARY) {
                moveArray("a.c[" + C + "].ids", "o.c[" + C + "].ids", 4);
                continue;
            }
            const int w = cs.encoding == PGPU_ENC_VARCELL ? 16 : t.isVarWidth() ? 0 : t.width();
            if (w != 1 && w != 4 && w != 8 && w != 16) { sweepOk = false; break; }
            moveArray("a.c[" + C + "].values", "o.c[" + C + "].values", w);
            if (cs.nullable) moveArray("a.c[" + C + "].nulls", "o.c[" + C + "].nulls", 1);
        }
        if (sweepOk) {
            src << "    PGPU_DEV static void move_tile(const PageArgs<NC>& a, const PageArgs<NC>& o, i64 tbase, int nt, u32 total, const unsigned short* sidx,\n"
                << "                                   const unsigned short* spid, const u64* gbase, const u32* lstart, unsigned char* stage) {\n"
                << mv.str() << "    }\n";
            res.sweepOk = true;
            res.sweepRowBytes = sweepRowBytes;
            res.sweepStageWidth = sweepStageWidth;
        }
    }
    src << "};\n\n";
    if (res.mode == KernelMode::AGG_GLOBAL) {
        src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_page(PageArgs<P::NC> a, Scratch sc) { body_agg_global<P>(a, sc); }\n";
        src << "extern \"C\" __global__ void pgpu_k_emit(Scratch sc, OutArgs<P::NO> out) { body_emit_global<P>(sc, out); }\n";
        if (finalStep)  // gather exchange: FINAL over the partial pages in this GPU's inbox
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_gather(const PageArgs<P::NC>* pages, GatherArgs g, Scratch sc) { body_gather_global<P>(pages, g, sc); }\n";
    } else {
        src << "extern \"C\" __global__ void __launch_bounds__(256, " << (getenv("PGPU_TUNE_MINBLOCKS") ? atoi(getenv("PGPU_TUNE_MINBLOCKS")) : (res.useDir ? 2 : 3)) << ") pgpu_k_page(PageArgs<P::NC> a, TableArgs t, Scratch sc) { body_agg_hash<P>(a, t, sc); }\n";
        if (finalStep)
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_gather(const PageArgs<P::NC>* pages, GatherArgs g, TableArgs t, Scratch sc) { body_gather_hash<P>(pages, g, t, sc); }\n";
        if (res.staged)
            src << "extern \"C\" __global__ void __launch_bounds__(512, 1) pgpu_k_page_staged(PageArgs<P::NC> a, TableArgs t, Scratch sc, int nstages) { body_agg_hash_staged<P>(a, t, sc, nstages); }\n";
        if (res.buckets) {
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_emit(TableArgs t, OutArgs<P::NO> out, Scratch sc, i64 first_slot, i64 nslots) { body_materialize_b<P>(t, out, sc, first_slot, nslots); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_rehash(TableArgs from, TableArgs to, Scratch sc) { body_rehash_b<P>(from, to, sc); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_count(TableArgs t, Scratch sc) { body_count_groups<P>(t, sc); }\n";
        } else {
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_emit(TableArgs t, OutArgs<P::NO> out, Scratch sc, i64 first_slot, i64 nslots) { body_materialize<P>(t, out, sc, first_slot, nslots); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_rehash(TableArgs from, TableArgs to, Scratch sc) { body_rehash<P>(from, to, sc); }\n";
            if (res.rangeOk) {
                src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_emit_range(TableArgs t, OutArgs<P::NO> out, Scratch sc, i64 first_slot, i64 nslots, RangeArgs rg) { body_materialize_range<P>(t, out, sc, first_slot, nslots, rg); }\n";
                src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_keyrange(TableArgs t, i64* out) { body_key_range<P>(t, out); }\n";
                src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_emit_probe(TableArgs t, OutArgs<P::NO> out, Scratch sc, ProbeArgs pa) { body_emit_probe<P>(t, out, sc, pa); }\n";
            }
        }
        // streaming runs (device header, "streaming runs"): SINGLE step over a one-key-per-slot tagless table
        res.streamOk = highCard && !res.buckets && plan.step == Step::SINGLE && plan.maxPartialGroups == 0 && envInt("PGPU_TUNE_STREAM", 1) != 0;
        if (res.streamOk) {
            // registers: 4 rows x (key + hot cells) stay live through the scan
            res.streamSub = std::min(std::max(envInt("PGPU_TUNE_STREAM_SUB", 1), 1), 4);
            const int streamMinBlocks = envInt("PGPU_TUNE_STREAM_MINBLOCKS", (KW + NHA) * res.streamSub <= 4 ? 4 : (KW + NHA) * res.streamSub <= 8 ? 3 : 2);
            src << "extern \"C\" __global__ void __launch_bounds__(256, " << streamMinBlocks << ") pgpu_k_runs_count(PageArgs<P::NC> a, u32* counts, Scratch sc) { body_runs_count<P>(a, counts, sc); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256, " << streamMinBlocks << ") pgpu_k_runs_write(PageArgs<P::NC> a, const u64* offsets, RunArgs ra, Scratch sc) { body_runs_write<P>(a, offsets, ra, sc); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_runs_emit(const u64* buf, i64 nrows, OutArgs<P::NO> out, Scratch sc, u32* violations) { body_runs_emit<P>(buf, nrows, out, sc, violations); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_runs_to_table(const u64* buf, u64 nrows, TableArgs to, Scratch sc) { body_runs_to_table<P>(buf, nrows, to, sc); }\n";
        }
        if (res.sweepOk) {
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_part_count(PageArgs<P::NC> a, PartArgs pa, Scratch sc) { body_part_count<P>(a, pa, sc); }\n";
            src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_part_scatter(PageArgs<P::NC> a, PageArgs<P::NC> out, PartArgs pa) { body_part_scatter<P>(a, out, pa); }\n";
            const int sweepMinBlocks = std::min(std::max(envInt("PGPU_TUNE_SWEEP_MINBLOCKS", 4), 1), 8);
            src << "extern \"C\" __global__ void __launch_bounds__(256, " << sweepMinBlocks << ") pgpu_k_page_sweep(PageArgs<P::NC> a, TableArgs t, Scratch sc, SweepArgs sw) { body_agg_hash_sweep<P>(a, t, sc, sw); }\n";
        }
        src << "extern \"C\" __global__ void __launch_bounds__(256) pgpu_k_replay(PageArgs<P::NC> a, TableArgs t, Scratch sc, const u64* rows, u64 n) { body_replay<P>(a, t, sc, rows, n); }\n";
    }
    res.source = src.str();
    if (res.streamOk && res.streamSub != 1) {  // experiment knob: rows per thread of the streaming-runs kernels = 4 * streamSub
        const std::string inc = "#include \"pgpu_device.cuh\"";
        res.source.insert(res.source.find(inc), "#define PGPU_STREAM_SUB " + std::to_string(res.streamSub) + "\n");
    }
    if (res.sweepOk && envInt("PGPU_TUNE_SWEEP_WT", 1024) != 1024) {  // experiment knob: rows per warp ticket of the sweep kernel (128 .. 4096, a multiple of 128)
        const int wt = std::min(std::max(envInt("PGPU_TUNE_SWEEP_WT", 1024) / 128 * 128, 128), 4096);
        const std::string inc = "#include \"pgpu_device.cuh\"";
        res.source.insert(res.source.find(inc), "#define PGPU_SWEEP_WT " + std::to_string(wt) + "\n");
    }
    if (res.sweepOk && envInt("PGPU_TUNE_SWEEP_INFLIGHT", 2) != 2) {  // experiment knob: runs per lane in flight in the sweep kernel (1 .. 4)
        const int inf = std::min(std::max(envInt("PGPU_TUNE_SWEEP_INFLIGHT", 2), 1), 4);
        const std::string inc = "#include \"pgpu_device.cuh\"";
        res.source.insert(res.source.find(inc), "#define PGPU_SWEEP_INFLIGHT " + std::to_string(inf) + "\n");
    }
    if (locB > 0) {
        const std::string inc = "#include \"pgpu_device.cuh\"";
        const size_t at = res.source.find(inc);
        res.source.insert(at, "#define PGPU_LOCB " + std::to_string(locB) + "\n#define PGPU_LOCSH " + std::to_string(locSh) + "\n");
    }
    return res;
}

}  // namespace pgpu
