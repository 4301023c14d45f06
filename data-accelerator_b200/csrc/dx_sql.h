// dx_sql.h — parser for the Spark-SQL subset that Data Accelerator flows actually contain
// (SURVEY.md §7.4): SELECT [DISTINCT] ... FROM t [[INNER | LEFT [OUTER]] JOIN u ON ...] [WHERE ...] [GROUP BY ...] [HAVING ...]
// [UNION [ALL] SELECT ...], expressions with AND/OR/NOT, comparisons, IN, BETWEEN, LIKE, IS NULL,
// arithmetic, function calls, CAST, dotted and back-ticked names.
//
// The statements reaching this parser are what TransformSQLParser hands to spark.sql(...)
// (reference: datax/processor/CommonProcessorFactory.scala:249-294) and the selectExpr strings of
// the projection file (CommonProcessorFactory.scala:99-100).  Anything outside the subset is a
// parse/plan error at flow-create time; there is no fallback engine.
#pragma once
#include <string>
#include <vector>
#include <memory>
#include <stdexcept>
#include <algorithm>

namespace dxsql {

struct SqlError : std::runtime_error { using std::runtime_error::runtime_error; };

enum class EK { Col, Int, Dec, Str, Null, Bool, Bin, Not, Neg, IsNull, In, Between, Like, Func, Cast, Star };

struct Expr;
typedef std::shared_ptr<Expr> ExprP;
struct Expr {
    EK kind;
    std::vector<std::string> parts;   // Col: name parts; Star: optional qualifier
    long long ival = 0;               // Int / Bool
    double dval = 0;                  // Dec
    std::string sval;                 // Str literal, Bin operator, Func name (upper-cased), Cast type (lower-cased)
    std::vector<ExprP> args;          // operands
    bool negated = false;             // IsNull / In / Between / Like
    bool distinct = false;            // COUNT(DISTINCT x)
    bool star_arg = false;            // COUNT(*)
    std::string alias;                // STRUCT(x AS name): alias on a function argument
    int depth = 1;                    // height of the tree below (the parser refuses more than DX_SQL_MAX_DEPTH: its consumers recurse)
};
#define DX_SQL_MAX_DEPTH 1000
#define DX_SQL_MAX_NESTING 256

struct SelectItem { ExprP expr; std::string alias; };
struct JoinClause { std::string table, alias; ExprP on; bool left_outer = false; };
struct SelectCore {
    bool distinct = false;
    std::vector<SelectItem> items;
    std::string from, from_alias;
    std::vector<JoinClause> joins;
    ExprP where;
    std::vector<ExprP> group_by;
    ExprP having;
};
struct Statement {
    std::vector<SelectCore> selects;     // > 1: UNION [ALL] chain
    std::vector<bool> union_all;         // between selects[i] and selects[i+1]
};

static inline std::string upper(std::string s) { for (auto& c : s) c = (char)toupper((unsigned char)c); return s; }
static inline std::string lower(std::string s) { for (auto& c : s) c = (char)tolower((unsigned char)c); return s; }

struct Tok { enum T { Id, QId, Num, Str, Op, End } t; std::string s; };

class Lexer {
    const std::string& src; size_t i = 0;
public:
    explicit Lexer(const std::string& s) : src(s) {}
    Tok next() {
        while (i < src.size() && isspace((unsigned char)src[i])) i++;
        if (i >= src.size()) return {Tok::End, ""};
        char c = src[i];
        if (isalpha((unsigned char)c) || c == '_') {
            size_t j = i; while (j < src.size() && (isalnum((unsigned char)src[j]) || src[j] == '_')) j++;
            Tok t{Tok::Id, src.substr(i, j - i)}; i = j; return t;
        }
        if (c == '`') {
            size_t j = src.find('`', i + 1);
            if (j == std::string::npos) throw SqlError("unterminated back-tick name");
            Tok t{Tok::QId, src.substr(i + 1, j - i - 1)}; i = j + 1; return t;
        }
        if (isdigit((unsigned char)c) || (c == '.' && i + 1 < src.size() && isdigit((unsigned char)src[i + 1]))) {
            size_t j = i; bool dot = false, ex = false;
            while (j < src.size()) {
                char d = src[j];
                if (isdigit((unsigned char)d)) j++;
                else if (d == '.' && !dot && !ex) { dot = true; j++; }
                else if ((d == 'e' || d == 'E') && !ex && j + 1 < src.size() &&
                         (isdigit((unsigned char)src[j + 1]) || ((src[j + 1] == '+' || src[j + 1] == '-') && j + 2 < src.size() && isdigit((unsigned char)src[j + 2])))) { ex = true; j += 2; }
                else break;
            }
            Tok t{Tok::Num, src.substr(i, j - i)}; i = j; return t;
        }
        if (c == '\'' || c == '"') {
            std::string out; size_t j = i + 1;
            while (true) {
                if (j >= src.size()) throw SqlError("unterminated string literal");
                char d = src[j++];
                if (d == c) break;
                if (d == '\\' && j < src.size()) {
                    char e = src[j++];
                    switch (e) { case 'n': out.push_back('\n'); break; case 't': out.push_back('\t'); break;
                                 case 'r': out.push_back('\r'); break; case '0': out.push_back('\0'); break;
                                 case '%': out += "\\%"; break; case '_': out += "\\_"; break;
                                 default: out.push_back(e); }
                } else out.push_back(d);
            }
            i = j; return {Tok::Str, out};
        }
        static const char* two[] = {"<>", "!=", "<=", ">=", "==", "||"};
        for (auto t : two) if (src.compare(i, 2, t) == 0) { i += 2; return {Tok::Op, t}; }
        i++;
        return {Tok::Op, std::string(1, c)};
    }
};

class Parser {
    std::vector<Tok> toks; size_t p = 0;
    const Tok& peek(size_t k = 0) const { return toks[std::min(p + k, toks.size() - 1)]; }
    Tok take() { Tok t = peek(); if (p < toks.size() - 1) p++; return t; }
    bool is_kw(const char* kw, size_t k = 0) const { return peek(k).t == Tok::Id && upper(peek(k).s) == kw; }
    bool is_op(const char* op, size_t k = 0) const { return peek(k).t == Tok::Op && peek(k).s == op; }
    bool accept_kw(const char* kw) { if (is_kw(kw)) { take(); return true; } return false; }
    bool accept_op(const char* op) { if (is_op(op)) { take(); return true; } return false; }
    void expect_op(const char* op) { if (!accept_op(op)) throw SqlError(std::string("expected '") + op + "' near '" + peek().s + "'"); }
    void expect_kw(const char* kw) { if (!accept_kw(kw)) throw SqlError(std::string("expected ") + kw + " near '" + peek().s + "'"); }
    static bool reserved(const std::string& u) {
        static const char* r[] = {"SELECT", "FROM", "WHERE", "GROUP", "BY", "AS", "AND", "OR", "NOT", "IN", "BETWEEN", "LIKE", "IS",
                                  "JOIN", "INNER", "ON", "UNION", "ALL", "DISTINCT", "HAVING", "ORDER", "LIMIT", "LEFT", "RIGHT",
                                  "FULL", "OUTER", "CROSS", "CASE", "WHEN", "THEN", "ELSE", "END"};
        for (auto k : r) if (u == k) return true;
        return false;
    }
    static ExprP mk(EK k) { auto e = std::make_shared<Expr>(); e->kind = k; return e; }
    // called once the operands are in place
    static ExprP fin(ExprP e) {
        int d = 0; for (auto& a : e->args) if (a && a->depth > d) d = a->depth;
        e->depth = d + 1;
        if (e->depth > DX_SQL_MAX_DEPTH) throw SqlError("expression deeper than " + std::to_string(DX_SQL_MAX_DEPTH) + " levels");
        return e;
    }
    int nesting = 0;
    struct Nest { int& n; explicit Nest(int& x) : n(x) { if (++n > DX_SQL_MAX_NESTING) { --n; throw SqlError("expression nested deeper than " + std::to_string(DX_SQL_MAX_NESTING) + " levels"); } } ~Nest() { --n; } };

    ExprP primary() {
        Nest nest(nesting);
        const Tok& t = peek();
        if (t.t == Tok::Num) {
            take();
            if (t.s.find_first_of(".eE") == std::string::npos) {
                auto e = mk(EK::Int);
                try { e->ival = std::stoll(t.s); } catch (...) { auto d = mk(EK::Dec); d->dval = std::stod(t.s); d->sval = t.s; return d; }
                return e;
            }
            auto e = mk(EK::Dec); e->dval = std::stod(t.s); e->sval = t.s; return e;
        }
        if (t.t == Tok::Str) { take(); auto e = mk(EK::Str); e->sval = t.s; return e; }
        if (t.t == Tok::Op && t.s == "(") { take(); ExprP e = expr(); expect_op(")"); return e; }
        if (t.t == Tok::Op && t.s == "*") { take(); return mk(EK::Star); }
        if (t.t == Tok::Op && t.s == "-") { take(); auto e = mk(EK::Neg); e->args.push_back(unary()); return fin(e); }
        if (t.t == Tok::Op && t.s == "+") { take(); return unary(); }
        if (t.t == Tok::Id || t.t == Tok::QId) {
            std::string u = upper(t.s);
            if (t.t == Tok::Id) {
                if (u == "NULL") { take(); return mk(EK::Null); }
                if (u == "TRUE" || u == "FALSE") { take(); auto e = mk(EK::Bool); e->ival = (u == "TRUE"); return e; }
                if (u == "CASE") {
                    // CASE [x] WHEN a THEN v ... [ELSE w] END  ==  IF(cond1, v1, IF(cond2, v2, ... w | NULL)): same evaluation order and
                    // the same treatment of a NULL condition (not taken) as Spark's CaseWhen
                    take();
                    ExprP subject;
                    if (!is_kw("WHEN")) subject = expr();
                    std::vector<std::pair<ExprP, ExprP>> arms;
                    while (accept_kw("WHEN")) {
                        ExprP c = expr();
                        if (subject) { auto eq = mk(EK::Bin); eq->sval = "="; eq->args = {subject, c}; c = fin(eq); }
                        expect_kw("THEN");
                        arms.emplace_back(c, expr());
                    }
                    if (arms.empty()) throw SqlError("CASE without WHEN");
                    ExprP tail = accept_kw("ELSE") ? expr() : mk(EK::Null);
                    expect_kw("END");
                    for (size_t i = arms.size(); i-- > 0;) { auto f = mk(EK::Func); f->sval = "IF"; f->args = {arms[i].first, arms[i].second, tail}; tail = fin(f); }
                    return tail;
                }
                if (u == "CAST" && is_op("(", 1)) {
                    take(); take();
                    auto e = mk(EK::Cast); e->args.push_back(expr());
                    expect_kw("AS");
                    Tok ty = take(); if (ty.t != Tok::Id) throw SqlError("expected type name in CAST");
                    e->sval = lower(ty.s);
                    if (accept_op("(")) { while (!is_op(")") && peek().t != Tok::End) take(); expect_op(")"); }
                    expect_op(")");
                    return fin(e);
                }
                if (is_op("(", 1) && !reserved(u)) {
                    take(); take();
                    auto e = mk(EK::Func); e->sval = u;
                    if (accept_kw("DISTINCT")) e->distinct = true;
                    if (!is_op(")")) {
                        if (is_op("*") && is_op(")", 1)) { take(); e->star_arg = true; }
                        else { e->args.push_back(arg()); while (accept_op(",")) e->args.push_back(arg()); }
                    }
                    expect_op(")");
                    return fin(e);
                }
                if (reserved(u)) throw SqlError("unexpected keyword '" + t.s + "'");
            }
            auto e = mk(EK::Col);
            e->parts.push_back(take().s);
            while (is_op(".")) {
                if (is_op("*", 1)) { take(); take(); e->kind = EK::Star; return e; }
                if (peek(1).t != Tok::Id && peek(1).t != Tok::QId) break;
                take(); e->parts.push_back(take().s);
            }
            return e;
        }
        throw SqlError("unexpected token '" + t.s + "'");
    }
    ExprP arg() {   // function argument, optionally aliased as in STRUCT(a AS b)
        ExprP e = expr();
        if (is_kw("AS")) {
            take(); Tok t = take(); if (t.t != Tok::Id && t.t != Tok::QId) throw SqlError("expected alias");
            auto c = std::make_shared<Expr>(*e); c->alias = t.s; return c;
        }
        return e;
    }
    ExprP unary() { return primary(); }
    ExprP mul() {
        ExprP l = unary();
        while (is_op("*") || is_op("/") || is_op("%")) { auto e = mk(EK::Bin); e->sval = take().s; e->args = {l, unary()}; l = fin(e); }
        return l;
    }
    ExprP add() {
        ExprP l = mul();
        while (is_op("+") || is_op("-") || is_op("||")) { auto e = mk(EK::Bin); e->sval = take().s; e->args = {l, mul()}; l = fin(e); }
        return l;
    }
    ExprP cmp() {
        ExprP l = add();
        for (;;) {
            if (is_op("=") || is_op("==") || is_op("<>") || is_op("!=") || is_op("<") || is_op("<=") || is_op(">") || is_op(">=")) {
                std::string op = take().s; if (op == "==") op = "="; if (op == "!=") op = "<>";
                auto e = mk(EK::Bin); e->sval = op; e->args = {l, add()}; l = fin(e); continue;
            }
            bool neg = false; size_t save = p;
            if (is_kw("NOT") && (is_kw("IN", 1) || is_kw("BETWEEN", 1) || is_kw("LIKE", 1))) { take(); neg = true; }
            if (accept_kw("IN")) {
                expect_op("(");
                auto e = mk(EK::In); e->negated = neg; e->args.push_back(l);
                e->args.push_back(expr()); while (accept_op(",")) e->args.push_back(expr());
                expect_op(")"); l = fin(e); continue;
            }
            if (accept_kw("BETWEEN")) {
                auto e = mk(EK::Between); e->negated = neg; e->args.push_back(l);
                e->args.push_back(add()); expect_kw("AND"); e->args.push_back(add()); l = fin(e); continue;
            }
            if (accept_kw("LIKE")) { auto e = mk(EK::Like); e->negated = neg; e->args = {l, add()}; l = fin(e); continue; }
            if (is_kw("IS")) {
                take(); bool n = accept_kw("NOT"); expect_kw("NULL");
                auto e = mk(EK::IsNull); e->negated = n; e->args.push_back(l); l = fin(e); continue;
            }
            p = save;
            return l;
        }
    }
    ExprP not_() {
        if (is_kw("NOT")) { Nest nest(nesting); take(); auto e = mk(EK::Not); e->args.push_back(not_()); return fin(e); }
        return cmp();
    }
    ExprP and_() {
        ExprP l = not_();
        while (is_kw("AND")) { take(); auto e = mk(EK::Bin); e->sval = "AND"; e->args = {l, not_()}; l = fin(e); }
        return l;
    }
    ExprP or_() {
        ExprP l = and_();
        while (is_kw("OR")) { take(); auto e = mk(EK::Bin); e->sval = "OR"; e->args = {l, and_()}; l = fin(e); }
        return l;
    }
    std::string alias_opt() {
        if (accept_kw("AS")) { Tok t = take(); if (t.t != Tok::Id && t.t != Tok::QId) throw SqlError("expected alias"); return t.s; }
        if ((peek().t == Tok::Id && !reserved(upper(peek().s))) || peek().t == Tok::QId) return take().s;
        return "";
    }
    SelectCore select_core() {
        SelectCore s;
        expect_kw("SELECT");
        if (accept_kw("DISTINCT")) s.distinct = true; else accept_kw("ALL");
        do { SelectItem it; it.expr = expr(); it.alias = alias_opt(); s.items.push_back(it); } while (accept_op(","));
        expect_kw("FROM");
        Tok t = take(); if (t.t != Tok::Id && t.t != Tok::QId) throw SqlError("expected table name after FROM");
        s.from = t.s; s.from_alias = alias_opt();
        for (;;) {
            if (is_kw("RIGHT") || is_kw("FULL") || is_kw("CROSS"))
                throw SqlError("only INNER and LEFT [OUTER] JOIN are inside the supported SQL subset");
            bool left_outer = false;
            if (is_kw("LEFT") && (is_kw("JOIN", 1) || (is_kw("OUTER", 1) && is_kw("JOIN", 2)))) { take(); accept_kw("OUTER"); left_outer = true; }
            else if (is_kw("INNER") && is_kw("JOIN", 1)) take();
            if (!accept_kw("JOIN")) break;
            JoinClause j; j.left_outer = left_outer; Tok jt = take(); if (jt.t != Tok::Id && jt.t != Tok::QId) throw SqlError("expected table name after JOIN");
            j.table = jt.s; j.alias = alias_opt();
            expect_kw("ON"); j.on = expr();
            s.joins.push_back(j);
        }
        if (accept_kw("WHERE")) s.where = expr();
        if (is_kw("GROUP")) { take(); expect_kw("BY"); do { s.group_by.push_back(expr()); } while (accept_op(",")); }
        if (accept_kw("HAVING")) s.having = expr();
        if (is_kw("ORDER") || is_kw("LIMIT"))
            throw SqlError(upper(peek().s) + " is outside the supported SQL subset");
        return s;
    }
public:
    explicit Parser(const std::string& text) {
        Lexer lx(text);
        for (;;) { Tok t = lx.next(); toks.push_back(t); if (t.t == Tok::End) break; }
    }
    ExprP expr() { return or_(); }
    // one selectExpr string: expression with optional alias
    SelectItem select_item() {
        SelectItem it; it.expr = expr(); it.alias = alias_opt();
        if (peek().t != Tok::End) throw SqlError("trailing input near '" + peek().s + "'");
        return it;
    }
    ExprP condition() { ExprP e = expr(); if (peek().t != Tok::End) throw SqlError("trailing input near '" + peek().s + "'"); return e; }
    Statement statement() {
        Statement st;
        st.selects.push_back(select_core());
        while (is_kw("UNION")) {
            take(); bool all = accept_kw("ALL"); if (!all) accept_kw("DISTINCT");
            st.union_all.push_back(all);
            st.selects.push_back(select_core());
        }
        accept_op(";");
        if (peek().t != Tok::End) throw SqlError("trailing input near '" + peek().s + "'");
        return st;
    }
};

// canonical text form (tests pin the parser through this)
static inline std::string to_string(const ExprP& e);
static inline std::string join_args(const std::vector<ExprP>& a, size_t from = 0) {
    std::string s; for (size_t i = from; i < a.size(); i++) { if (i > from) s += ", "; s += to_string(a[i]); } return s;
}
static inline std::string to_string_noalias(const ExprP& e);
static inline std::string to_string(const ExprP& e) {
    if (!e) return "<null>";
    std::string s = to_string_noalias(e);
    return e->alias.empty() ? s : s + " AS " + e->alias;
}
static inline std::string to_string_noalias(const ExprP& e) {
    switch (e->kind) {
    case EK::Col: { std::string s; for (size_t i = 0; i < e->parts.size(); i++) { if (i) s += "."; s += e->parts[i]; } return s; }
    case EK::Int: return std::to_string(e->ival);
    case EK::Dec: return e->sval;
    case EK::Str: return "'" + e->sval + "'";
    case EK::Null: return "NULL";
    case EK::Bool: return e->ival ? "TRUE" : "FALSE";
    case EK::Bin: return "(" + to_string(e->args[0]) + " " + e->sval + " " + to_string(e->args[1]) + ")";
    case EK::Not: return "(NOT " + to_string(e->args[0]) + ")";
    case EK::Neg: return "(-" + to_string(e->args[0]) + ")";
    case EK::IsNull: return "(" + to_string(e->args[0]) + (e->negated ? " IS NOT NULL)" : " IS NULL)");
    case EK::In: return "(" + to_string(e->args[0]) + (e->negated ? " NOT IN (" : " IN (") + join_args(e->args, 1) + "))";
    case EK::Between: return "(" + to_string(e->args[0]) + (e->negated ? " NOT BETWEEN " : " BETWEEN ") + to_string(e->args[1]) + " AND " + to_string(e->args[2]) + ")";
    case EK::Like: return "(" + to_string(e->args[0]) + (e->negated ? " NOT LIKE " : " LIKE ") + to_string(e->args[1]) + ")";
    case EK::Func: return e->sval + "(" + (e->distinct ? "DISTINCT " : "") + (e->star_arg ? "*" : join_args(e->args)) + ")";
    case EK::Cast: return "CAST(" + to_string(e->args[0]) + " AS " + e->sval + ")";
    case EK::Star: { std::string s; for (auto& p : e->parts) s += p + "."; return s + "*"; }
    }
    return "?";
}
static inline std::string to_string(const Statement& st) {
    std::string out;
    for (size_t k = 0; k < st.selects.size(); k++) {
        const SelectCore& s = st.selects[k];
        if (k) out += st.union_all[k - 1] ? " UNION ALL " : " UNION ";
        out += s.distinct ? "SELECT DISTINCT " : "SELECT ";
        for (size_t i = 0; i < s.items.size(); i++) {
            if (i) out += ", ";
            out += to_string(s.items[i].expr);
            if (!s.items[i].alias.empty()) out += " AS " + s.items[i].alias;
        }
        out += " FROM " + s.from; if (!s.from_alias.empty()) out += " AS " + s.from_alias;
        for (auto& j : s.joins) { out += (j.left_outer ? " LEFT JOIN " : " JOIN ") + j.table; if (!j.alias.empty()) out += " AS " + j.alias; out += " ON " + to_string(j.on); }
        if (s.where) out += " WHERE " + to_string(s.where);
        if (!s.group_by.empty()) { out += " GROUP BY "; out += join_args(s.group_by); }
        if (s.having) out += " HAVING " + to_string(s.having);
    }
    return out;
}

} // namespace dxsql
