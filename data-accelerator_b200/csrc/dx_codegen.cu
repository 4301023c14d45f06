// dx_codegen.cu — the rules / flow code generator restated: user code with ProcessRules / ProcessAlerts /
// ProcessAggregateRules / ProcessAggregateAlerts / CreateMetric calls, OUTPUT ... TO, CREATE TABLE, WITH UPSERT and TIMEWINDOW(...)
// plus the rule definitions and the query / output templates  ->  the `--DataXQuery--` transform text the processor is handed.
//
// Restates DataX.Flow.CodegenRules: Engine.GenerateCode and its passes (Engine.cs:32-107 order of the passes, :142-198 auto-insertion
// of alert calls, :200-354 the four Process* passes, :356-378 CreateMetric, :380-397 the rule array, :399-493 template expansion per
// rule, :495-626 OUTPUT / CREATE TABLE / TIMEWINDOW / UPSERT / clean-up), the Rule model's name mangling (Rule.cs:76-281) and the SQL
// pretty-printer the result goes through (Formatter.cs: a token-driven indenter in the style of Hibernate's BasicFormatterImpl, with
// MAP / STRUCT / IF / ARRAY / filterNull opening a new indent level).  Validated byte for byte against the 18 expected outputs of
// CodegenTests.cs (tests/test_codegen_restatement.py).  Host code, design time: nothing here runs per micro-batch.
//
// .NET behaviours that matter and are reproduced: Regex `.` matches every character but '\n' (so also '\r'); a MatchCollection is
// taken over the text as it was when Matches() was called, while the replacements go into the changing text; String.Replace replaces
// every occurrence, a null replacement is the empty string; Trim() without arguments trims white space, TrimEnd(chars) only those.
#include "dx_common.h"
#include "dx_json.h"
#include "dx_codegen.h"
#include <regex>
#include <set>

namespace {
typedef std::string S;

S replace_all(S s, const S& from, const S& to) {
    if (from.empty()) return s;
    size_t at = 0;
    while ((at = s.find(from, at)) != S::npos) { s.replace(at, from.size(), to); at += to.size(); }
    return s;
}
bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v'; }
S trim(const S& s) { size_t a = 0, b = s.size(); while (a < b && is_ws(s[a])) a++; while (b > a && is_ws(s[b - 1])) b--; return s.substr(a, b - a); }
S trim_chars(const S& s, const char* chars) { size_t a = 0, b = s.size(); while (a < b && strchr(chars, s[a])) a++; while (b > a && strchr(chars, s[b - 1])) b--; return s.substr(a, b - a); }
S trim_end(const S& s, const char* chars) { size_t b = s.size(); while (b > 0 && strchr(chars, s[b - 1])) b--; return s.substr(0, b); }
S lower(S s) { for (auto& c : s) c = (char)tolower((unsigned char)c); return s; }
bool ends_with(const S& s, const S& t) { return s.size() >= t.size() && s.compare(s.size() - t.size(), t.size(), t) == 0; }
std::vector<S> split(const S& s, char c) { std::vector<S> o; size_t a = 0; for (;;) { size_t b = s.find(c, a); if (b == S::npos) { o.push_back(s.substr(a)); return o; } o.push_back(s.substr(a, b - a)); a = b + 1; } }
// a .NET pattern as an ECMAScript one: `.` excludes '\n' only
std::regex rx(const S& pattern, bool icase = true) {
    S p;
    for (size_t i = 0; i < pattern.size(); i++) {
        if (pattern[i] == '\\' && i + 1 < pattern.size()) { p += pattern[i]; p += pattern[++i]; continue; }
        if (pattern[i] == '[') { while (i < pattern.size() && pattern[i] != ']') p += pattern[i++]; p += ']'; continue; }
        if (pattern[i] == '.') p += "[^\\n]"; else p += pattern[i];
    }
    return std::regex(p, icase ? std::regex::ECMAScript | std::regex::icase : std::regex::ECMAScript);
}
// every match with its groups as strings of their own: the callers rewrite the text while they walk the matches (a .NET
// MatchCollection keeps the string it was made from)
typedef std::vector<S> Groups;
std::vector<Groups> matches(const S& text, const std::regex& r) {
    std::vector<Groups> out;
    for (auto it = std::sregex_iterator(text.begin(), text.end(), r); it != std::sregex_iterator(); ++it) {
        Groups g; for (size_t k = 0; k < it->size(); k++) g.push_back((*it)[k].str());
        out.push_back(g);
    }
    return out;
}

// ---- Rule (Rule.cs) ------------------------------------------------------------------------------------------------
struct Rule {
    S ruleId, productId, ruleType, ruleDescription, ruleCategory, severity, condition, tagname, tag, fact, id, outputTemplate, targetTable;
    std::vector<S> aggs, pivots, sinks, alertsinks;
    bool has_alertsinks = false, has_pivots = false, isalert = false;

    static void agg_parts(const S& agg, S* op, S* col) {           // "(.*)\((.*?)\)": greedy function part, lazy argument
        static const std::regex r = rx("(.*)\\((.*?)\\)", false);
        std::smatch m;
        if (!std::regex_search(agg, m, r)) throw std::runtime_error("aggregate '" + agg + "' is not of the form FUNC(column)");
        *op = m[1]; *col = m[2];
    }
    static S mangled(const S& agg) {                                // AVG(a.b) -> ab_AVG; min(`a.b`) -> `a.b_min`
        S op, col; agg_parts(agg, &op, &col);
        if (ends_with(col, "`")) return trim_end(col, "`") + "_" + op + "`";
        return replace_all(col, ".", "") + "_" + op;
    }
    static S list_to_string(const std::vector<S>& l) { S r; for (auto& s : l) r += s + ", "; return trim_end(r, " ,"); }
    static S last_section(const S& s) { if (s.find('.') == S::npos) return s; auto p = split(s, '.'); return p.back(); }
    S aggs_to_select() const {
        try { S r; for (auto& a : aggs) r += a + " AS " + mangled(a) + ", "; return trim_end(r, " ,"); } catch (...) { return ""; }
    }
    S condition_to_sql() const {
        if (aggs.empty()) return condition;
        try {
            S r = condition;
            for (auto& a : aggs) r = replace_all(r, a, mangled(a));
            for (auto& p : pivots) if (!(p.size() && p[0] == '`') && p.find('.') != S::npos) r = replace_all(r, p, last_section(p));
            return r;
        } catch (...) { return ""; }
    }
    S aggs_to_template() const {
        if (aggs.empty()) return "";
        try {
            std::vector<std::pair<S, std::vector<S>>> by_col;           // insertion-ordered, as a Dictionary enumerates
            for (auto& a : aggs) {
                S op, col; agg_parts(a, &op, &col);
                bool found = false;
                for (auto& kv : by_col) if (kv.first == col) { kv.second.push_back(op); found = true; }
                if (!found) by_col.push_back({col, {op}});
            }
            S r = "MAP(\n";
            for (auto& kv : by_col) {
                r += "'" + kv.first + "', MAP(\n";
                for (auto& op : kv.second) {
                    if (ends_with(kv.first, "`")) r += "'" + op + "', " + trim_end(kv.first, "`") + "_" + op + "`,";
                    else r += "'" + op + "', " + replace_all(kv.first, ".", "") + "_" + op + ",";
                }
                r = trim_end(r, " ,");
                r += "), \n";
            }
            r = trim_end(r, " ,\n");
            return r + "\n) AS aggs";
        } catch (...) { return ""; }
    }
    S pivots_to_template() const {
        S r;
        for (auto& p : pivots) {
            if (ends_with(trim(p), "`")) r += "'" + p + "', " + p + ",\n";
            else r += "'" + p + "', " + last_section(p) + ",\n";
        }
        return trim_end(r, " ,\n");
    }
    S rules_object() const {
        return "MAP('ruleId', '" + ruleId + "', 'ruleDescription', '" + ruleDescription + "', 'severity', '" + severity + "', '" + tagname + "', '" + tag + "')";
    }
};

// Json.NET binds a property by its exact name, else ignoring case ("$isAlert" fills $isalert)
const JVal* jget(const JVal& o, const char* k) {
    if (const JVal* v = o.get(k)) return v;
    for (auto& kv : o.obj) if (lower(kv.first) == lower(k)) return &kv.second;
    return nullptr;
}
S jstr(const JVal& o, const char* k) { const JVal* v = jget(o, k); return v && v->kind == JVal::STR ? v->str : S(); }
std::vector<S> jlist(const JVal& o, const char* k, bool* present) {
    std::vector<S> out; const JVal* v = jget(o, k);
    *present = v && v->kind == JVal::ARR;
    if (*present) for (auto& e : v->arr) out.push_back(e.kind == JVal::STR ? e.str : S());
    return out;
}
// Json.NET reads // and /* */ comments; the schema parser does not
S strip_json_comments(const S& j) {
    S o; bool in_str = false;
    for (size_t i = 0; i < j.size(); i++) {
        const char c = j[i];
        if (in_str) { o += c; if (c == '\\' && i + 1 < j.size()) o += j[++i]; else if (c == '"') in_str = false; continue; }
        if (c == '"') { in_str = true; o += c; continue; }
        if (c == '/' && i + 1 < j.size() && j[i + 1] == '/') { while (i < j.size() && j[i] != '\n') i++; o += '\n'; continue; }
        if (c == '/' && i + 1 < j.size() && j[i + 1] == '*') { size_t e = j.find("*/", i + 2); i = e == S::npos ? j.size() : e + 1; continue; }
        o += c;
    }
    return o;
}
std::vector<Rule> load_rules(const S& json) {
    JVal root = dx_parse_json(strip_json_comments(json));
    if (root.kind != JVal::ARR) throw std::runtime_error("the rules must be a JSON array");
    std::vector<Rule> out;
    for (auto& o : root.arr) {
        Rule r; bool p;
        r.ruleId = jstr(o, "$ruleId"); r.productId = jstr(o, "$productId"); r.ruleType = jstr(o, "$ruleType"); r.ruleDescription = jstr(o, "$ruleDescription");
        r.ruleCategory = jstr(o, "$ruleCategory"); r.severity = jstr(o, "$severity"); r.condition = jstr(o, "$condition"); r.tagname = jstr(o, "$tagname");
        r.tag = jstr(o, "$tag"); r.fact = jstr(o, "$fact"); r.id = jstr(o, "$id"); r.outputTemplate = jstr(o, "$outputTemplate"); r.targetTable = jstr(o, "schemaTableName");
        r.aggs = jlist(o, "$aggs", &p); r.pivots = jlist(o, "$pivots", &r.has_pivots); r.sinks = jlist(o, "$sinks", &p); r.alertsinks = jlist(o, "$alertsinks", &r.has_alertsinks);
        const JVal* a = jget(o, "$isalert"); r.isalert = a && a->kind == JVal::BOOL && a->b;
        out.push_back(r);
    }
    return out;
}

// ---- templates: <queries><query type="..">text</query>..</queries>, <outputTemplates><outputTemplate id="..">text</outputTemplate>..
S xml_text(S s) {
    S o;
    for (size_t i = 0; i < s.size();) {
        if (s.compare(i, 9, "<![CDATA[") == 0) { size_t e = s.find("]]>", i); if (e == S::npos) e = s.size(); o += s.substr(i + 9, e - i - 9); i = e + 3; continue; }
        if (s[i] == '&') {
            static const std::pair<const char*, char> ent[] = {{"&lt;", '<'}, {"&gt;", '>'}, {"&amp;", '&'}, {"&quot;", '"'}, {"&apos;", '\''}};
            bool hit = false;
            for (auto& e : ent) if (s.compare(i, strlen(e.first), e.first) == 0) { o += e.second; i += strlen(e.first); hit = true; break; }
            if (hit) continue;
        }
        if (s[i] == '\r') { i++; if (i >= s.size() || s[i] != '\n') o += '\n'; continue; }        // XML line-end normalisation
        o += s[i++];
    }
    return o;
}
std::vector<std::pair<S, S>> load_templates(const S& xml, const S& element, const S& attribute) {
    std::vector<std::pair<S, S>> out;
    const std::regex open("<" + element + "\\s+" + attribute + "\\s*=\\s*\"([^\"]*)\"\\s*>");
    for (auto it = std::sregex_iterator(xml.begin(), xml.end(), open); it != std::sregex_iterator(); ++it) {
        const size_t body = (size_t)it->position(0) + (size_t)it->length(0);
        const size_t close = xml.find("</" + element + ">", body);
        if (close == S::npos) throw std::runtime_error("template XML: <" + element + "> is not closed");
        out.push_back({(*it)[1], xml_text(xml.substr(body, close - body))});
    }
    return out;
}
const S* find_template(const std::vector<std::pair<S, S>>& v, const S& key) { for (auto& kv : v) if (kv.first == key) return &kv.second; return nullptr; }

// ---- the SQL pretty-printer (Formatter.cs) ---------------------------------------------------------------------------
struct Printer {
    std::vector<S> toks; size_t at = 0;
    S out, tok, lc, last;
    std::vector<bool> saved_after; std::vector<int> saved_parens;
    bool begin_before_end = false, after_between = false, after_list_kw = false, after_insert = false, after_on = false, line_start = true,
         command_ended = false, in_ctor = false;
    int indent = 1, in_function = 0, parens_since_select = 0;

    static bool in(const char* const* set, const S& s) { for (; *set; set++) if (s == *set) return true; return false; }
    static bool begin_clause(const S& s) { static const char* k[] = {"left", "right", "inner", "outer", "group", "order", nullptr}; return in(k, s); }
    static bool end_clause(const S& s) { static const char* k[] = {"where", "set", "having", "join", "from", "by", "into", "union", nullptr}; return in(k, s); }
    static bool logical(const S& s) { static const char* k[] = {"and", "or", "when", "else", "end", nullptr}; return in(k, s); }
    static bool quantifier(const S& s) { static const char* k[] = {"in", "all", "exists", "some", "any", nullptr}; return in(k, s); }
    static bool dml(const S& s) { static const char* k[] = {"insert", "update", "delete", nullptr}; return in(k, s); }
    static bool misc(const S& s) { static const char* k[] = {"select", "on", "struct", "map", "if", "array", "filternull", nullptr}; return in(k, s); }
    static bool white(const S& s) { return S(" \n\r\f\t").find(s) != S::npos; }
    static bool function_name(const S& t) {
        if (t.empty()) return false;
        const unsigned char b = (unsigned char)t[0];
        const bool ident = isalpha(b) || b == '$' || b == '_' || b == '"' || b >= 0x80;
        return ident && !logical(t) && !end_clause(t) && !quantifier(t) && !dml(t) && !misc(t);
    }
    explicit Printer(const S& sql) {
        const S delim = "()+*/-=<>'`\"[],; \n\r\f\t";
        for (size_t i = 0; i < sql.size();) {
            if (delim.find(sql[i]) != S::npos) { toks.push_back(S(1, sql[i])); i++; continue; }
            size_t e = sql.find_first_of(delim, i); if (e == S::npos) e = sql.size();
            toks.push_back(sql.substr(i, e - i)); i = e;
        }
    }
    void emit() { out += tok; }
    void newline() { out += "\n"; for (int i = 0; i < indent; i++) out += "    "; line_start = true; }
    void quoted(const S& q) { while (at < toks.size()) { const S t = toks[at++]; tok += t; if (t == q) break; } }
    void other() {
        emit();
        if (lc == "between") after_between = true;
        if (after_insert) { newline(); after_insert = false; }
        else { line_start = false; if (lc == "case") indent++; }
    }
    S run() {
        out = "\n    ";
        while (at < toks.size()) {
            tok = toks[at++]; lc = lower(tok);
            if (tok == "'") quoted("'"); else if (tok == "\"") quoted("\"");
            if (tok == ";") { emit(); indent = 1; command_ended = true; newline(); }
            else if (after_list_kw && tok == ",") { emit(); newline(); }
            else if (after_on && tok == ",") { emit(); indent--; newline(); after_on = false; after_list_kw = true; }
            else if (tok == "(") {
                if (command_ended) emit();
                else {
                    if (function_name(last) || in_function > 0) in_function++;
                    line_start = false;
                    emit();
                    if (in_function == 0) {
                        if (last == "struct" || last == "map" || last == "if" || last == "array" || last == "filternull") in_ctor = true;
                        if (!after_list_kw || in_ctor) { indent++; newline(); line_start = true; }
                    }
                    parens_since_select++;
                }
            } else if (tok == ")") {
                if (command_ended) emit();
                else {
                    parens_since_select--;
                    if (parens_since_select < 0) {
                        indent--;
                        if (saved_parens.empty()) throw std::runtime_error("formatter: ')' without a matching '(' (the reference's formatter fails on its empty stack here)");
                        parens_since_select = saved_parens.back(); saved_parens.pop_back();
                        after_list_kw = saved_after.back(); saved_after.pop_back();
                    }
                    if (in_function > 0) { in_function--; emit(); }
                    else { if (!after_list_kw || in_ctor) { indent--; newline(); } emit(); }
                    line_start = false;
                }
            } else if (begin_clause(lc)) {
                if (!begin_before_end) { if (after_on) { indent--; after_on = false; } indent--; newline(); }
                emit(); line_start = false; begin_before_end = true;
            } else if (end_clause(lc)) {
                if (!begin_before_end) { indent--; if (after_on) { indent--; after_on = false; } newline(); }
                emit();
                if (lc != "union") indent++;
                newline();
                begin_before_end = false;
                after_list_kw = lc == "by" || lc == "set" || lc == "from";
            } else if (lc == "select") {
                emit(); indent++; newline();
                saved_parens.push_back(parens_since_select); saved_after.push_back(after_list_kw);
                parens_since_select = 0; after_list_kw = true; command_ended = false;
            } else if (dml(lc)) {
                emit(); indent++; line_start = false;
                if (lc == "update") newline();
                if (lc == "insert") after_insert = true;
                command_ended = false;
            } else if (lc == "values") { indent--; newline(); emit(); indent++; newline(); }
            else if (lc == "on") { indent++; after_on = true; newline(); emit(); line_start = false; }
            else if (after_between && lc == "and") { other(); after_between = false; }
            else if (logical(lc)) { if (lc == "end") indent--; newline(); emit(); line_start = false; }
            else if (white(tok)) { if (!line_start) out += " "; }
            else other();
            if (!white(tok)) last = lc;
        }
        return out;
    }
};

// ---- Engine (Engine.cs) ----------------------------------------------------------------------------------------------
struct Engine {
    std::vector<Rule> rules;
    std::vector<std::pair<S, S>> queries, outputs_t;
    S code;
    int statement = 0, rule_counter = 1;
    const S def_target = "DataXProcessedInput";

    std::vector<const Rule*> pick(const S& product, const S& type, const S& table, bool alerts_only) const {
        std::vector<const Rule*> o;
        for (auto& r : rules) if ((product.empty() || r.productId == product) && r.ruleType == type && r.targetTable == table && (!alerts_only || r.isalert)) o.push_back(&r);
        return o;
    }
    // Engine.cs:399-493
    S expand(const std::vector<const Rule*>& rs, const S& tmpl, const S& input_table) {
        if (rs.empty()) return "";
        rule_counter = 1;
        S result;
        static const std::regex apply = rx("ApplyTemplate\\s{0,}\\(\\s{0,}(.*?)\\s{0,},\\s{0,}(.*?)\\s{0,}\\)");
        for (const Rule* rule : rs) {
            result += trim(tmpl);
            for (auto& m : matches(result, apply)) {
                const S* o = nullptr; const S which = m[2];
                if (which == "$outputTemplate") {
                    if (!rule->outputTemplate.empty()) o = find_template(outputs_t, rule->outputTemplate);
                    else if (lower(rule->ruleType).find("aggregate") != S::npos) o = find_template(outputs_t, "defaultAggOutputTemplate");
                } else o = find_template(outputs_t, which);
                if (!o) result = replace_all(result, m[0], "SELECT * FROM " + m[1]);
                else {
                    S t = replace_all(*o, "$aggstemplate", rule->aggs_to_template());
                    t = replace_all(t, "$pivotstemplate", rule->pivots_to_template());
                    result = replace_all(result, m[0], "SELECT " + t + " FROM " + m[1]);
                }
            }
            if (!rule->has_alertsinks || (rule->alertsinks.size() == 1 && rule->alertsinks[0] == "Metrics")) {
                result = replace_all(result, "OUTPUT aa3_$ruleCounter TO $alertsinks;", "");
                result = replace_all(result, "OUTPUT sa2_$ruleCounter TO $alertsinks;", "");
            } else {
                std::vector<S> sinks; for (auto& s : rule->alertsinks) if (s != "Metrics") sinks.push_back(s);
                result = replace_all(result, "$alertsinks", Rule::list_to_string(sinks));
            }
            const S counter = std::to_string(statement) + "_" + std::to_string(rule_counter);
            const std::pair<const char*, S> subs[] = {
                {"$productId", rule->productId}, {"$ruleId", rule->ruleId}, {"$ruleCounter", counter}, {"$ruleDescription", rule->ruleDescription},
                {"$ruleCategory", rule->ruleCategory}, {"$ruleType", rule->ruleType}, {"$severity", rule->severity}, {"$aggs", rule->aggs_to_select()},
                {"$condition", rule->condition_to_sql()}, {"$tagname", rule->tagname}, {"$tag", rule->tag}, {"$sinks", Rule::list_to_string(rule->sinks)},
                {"$ruleObject", rule->rules_object()}, {"$id", rule->id}, {"$fact", rule->fact}, {"DataXProcessedInput", input_table}};
            for (auto& kv : subs) result = replace_all(result, kv.first, kv.second);
            if (rule->pivots.empty()) { result = replace_all(result, "GROUP BY $pivots", ""); result = replace_all(result, "$pivots,", ""); }
            else result = replace_all(result, "$pivots", Rule::list_to_string(rule->pivots));
            rule_counter++;
        }
        return replace_all(result, "\n", "\r\n");
    }
    S array_conditions(const std::vector<const Rule*>& rs) const {
        if (rs.empty()) return "'NULL'";
        S r = "filterNull(Array(\n";
        for (const Rule* rule : rs) {
            r += "IF($condition, $ruleObject, NULL),\n";
            r = replace_all(r, "$condition", rule->condition);
            r = replace_all(r, "$ruleObject", rule->rules_object());
        }
        return trim_end(r, ", \n") + "\n))";
    }
    const S& query(const S& type) const { const S* q = find_template(queries, type); if (!q) throw std::runtime_error("no query template of type '" + type + "'"); return *q; }

    void auto_alert_calls(const S& product) {                       // Engine.cs:142-198
        std::vector<std::pair<S, std::vector<S>>> kinds;
        for (auto& r : rules) {
            if (!r.isalert || (!product.empty() && r.productId != product)) continue;
            bool seen = false;
            for (auto& kv : kinds) if (kv.first == r.targetTable) { seen = true; if (std::find(kv.second.begin(), kv.second.end(), r.ruleType) == kv.second.end()) kv.second.push_back(r.ruleType); }
            if (!seen) kinds.push_back({r.targetTable, {r.ruleType}});
        }
        for (auto& kv : kinds) for (auto& type : kv.second) {
            const bool simple = type == "SimpleRule";
            const std::regex r = rx(S(simple ? "ProcessAlerts" : "ProcessAggregateAlerts") + "\\s{0,}\\(\\s{0,}" + kv.first + "\\s{0,}\\)");
            if (matches(code, r).empty()) code += S("\n") + (simple ? "ProcessAlerts(" : "ProcessAggregateAlerts(") + kv.first + ");";
        }
    }
    void alerts_pass(const S& product, const char* call, const char* type, const char* tmpl) {           // Engine.cs:200-230, 268-297
        for (auto& m : matches(code, rx(S(call) + "\\s{0,}\\(\\s{0,}(.*?)\\s{0,}\\)"))) {
            statement++;
            S table = m[1]; if (table.empty()) table = def_target;
            code = replace_all(code, m[0], expand(pick(product, type, table, true), query(tmpl), table));
        }
    }
    S run(const S& product) {
        auto_alert_calls(product);
        alerts_pass(product, "ProcessAlerts", "SimpleRule", "SimpleAlert");
        for (auto& m : matches(code, rx("(.*?)\\s{0,}=\\s{0,}ProcessRules\\s{0,}\\(\\s{0,}(.*?)\\s{0,}\\)"))) {             // Engine.cs:232-266
            statement++;
            S table = m[2]; if (table.empty()) table = def_target;
            S s = replace_all(query("SimpleRule"), "$arrayConditions", array_conditions(pick(product, "SimpleRule", table, false)));
            s = replace_all(s, "\n", "\r\n"); s = replace_all(s, "$return", m[1]); s = replace_all(s, def_target, table);
            code = replace_all(code, m[0], s);
        }
        for (auto& m : matches(code, rx("(.*?)\\s{0,}=\\s{0,}ProcessAggregateRules\\s{0,}\\(\\s{0,}(.*?)\\s{0,}\\)"))) {    // Engine.cs:299-354
            statement++;
            S table = m[2]; if (table.empty()) table = def_target;
            S s = expand(pick(product, "AggregateRule", table, false), query("AggregateRule"), table);
            s += "\n\n--DataXQuery--\nar4_" + std::to_string(statement) + " = ";
            for (int i = 1; i < rule_counter; i++) s += "SELECT * FROM ar3_" + std::to_string(statement) + "_" + std::to_string(i) + (i == rule_counter - 1 ? "" : " UNION ");
            s += "\n\n--DataXQuery--\n$return = SELECT * FROM ar4_" + std::to_string(statement);
            s = replace_all(s, "\n", "\r\n"); s = replace_all(s, "$return", m[1]);
            code = replace_all(code, m[0], s);
        }
        alerts_pass(product, "ProcessAggregateAlerts", "AggregateRule", "AggregateAlert");
        for (auto& m : matches(code, rx("(.*?)\\s{0,}=\\s{0,}CreateMetric\\s{0,}\\(\\s{0,}(.*?)\\s{0,},\\s{0,}(.*?)\\s{0,}\\)"))) {       // Engine.cs:356-378
            const S name = m[1], from = m[2], metric = m[3];
            code = replace_all(code, m[0], "\n\n--DataXQuery--\n" + name + " = SELECT DISTINCT DATE_TRUNC('second', current_timestamp()) AS EventTime, '" + name +
                               "' AS MetricName, " + metric + " AS Metric, '" + product + "' AS Product, '' AS Pivot1 FROM " + from + " GROUP BY EventTime, MetricName, Metric, Product, Pivot1;");
        }
        return code;
    }
};
}  // namespace

// S450_PrepareTransformFile.Process (DataX.Config/ConfigGeneration/Processor/S450_PrepareTransformFile.cs:48-64) +
// RuleDefinitionGenerator.GenerateRuleDefinitions (DataX.Config/InternalService/RuleDefinitionGenerator.cs:15-33): the flow's GUI
// JSON -> user code (gui.process.queries joined by '\n') and rule definitions (gui.rules[].properties with $productId = the flow's
// name), then the generator.  The stored form of a flow writes "_S_" for a leading "$" of a property name
// (LocalDesignTimeStorage.cs:169-174, CosmosDBUtility.cs:108-116); both spellings are read.  A rule without a $ruleId gets a GUID in
// the reference; here it gets "<flow name>-rule-<position>" so that the output is reproducible.
static void json_write(const JVal& v, S* o) {
    switch (v.kind) {
    case JVal::NUL: *o += "null"; break;
    case JVal::BOOL: *o += v.b ? "true" : "false"; break;
    case JVal::NUM: { char b[64]; if (v.num == (double)(long long)v.num) snprintf(b, sizeof b, "%lld", (long long)v.num); else snprintf(b, sizeof b, "%.17g", v.num); *o += b; break; }
    case JVal::STR: { *o += '"'; for (unsigned char c : v.str) { if (c == '"' || c == '\\') { *o += '\\'; *o += (char)c; } else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); *o += b; } else *o += (char)c; } *o += '"'; break; }
    case JVal::ARR: { *o += '['; for (size_t i = 0; i < v.arr.size(); i++) { if (i) *o += ','; json_write(v.arr[i], o); } *o += ']'; break; }
    case JVal::OBJ: { *o += '{'; for (size_t i = 0; i < v.obj.size(); i++) { if (i) *o += ','; JVal k; k.kind = JVal::STR; k.str = v.obj[i].first; json_write(k, o); *o += ':'; json_write(v.obj[i].second, o); } *o += '}'; break; }
    }
}
int dx_codegen_flow_impl(const std::string& flow_json, const std::string& query_xml, const std::string& output_xml, dx_codegen_result* res, std::string* err) {
    try {
        S text = flow_json.compare(0, 3, "\xEF\xBB\xBF") == 0 ? flow_json.substr(3) : flow_json;
        text = replace_all(text, "\"_S_", "\"$");
        JVal root = dx_parse_json(strip_json_comments(text));
        const JVal* gui = root.get("gui"); if (!gui) gui = &root;
        const JVal* name = gui->get("name");
        if (!name || name->kind != JVal::STR) throw std::runtime_error("flow JSON: no name");
        S code;
        if (const JVal* pr = gui->get("process")) if (const JVal* q = pr->get("queries")) for (size_t i = 0; i < q->arr.size(); i++) code += (i ? "\n" : "") + q->arr[i].str;
        JVal defs; defs.kind = JVal::ARR;
        if (const JVal* rules = gui->get("rules")) for (size_t i = 0; i < rules->arr.size(); i++) {
            const JVal* props = rules->arr[i].get("properties"); if (!props) continue;
            JVal r = *props; bool have_id = false, have_product = false;
            for (auto& kv : r.obj) {
                if (kv.first == "$ruleId") { have_id = true; if (kv.second.kind != JVal::STR || kv.second.str.empty()) { kv.second.kind = JVal::STR; kv.second.str = name->str + "-rule-" + std::to_string(i + 1); } }
                if (kv.first == "$productId") { have_product = true; kv.second.kind = JVal::STR; kv.second.str = name->str; }
            }
            JVal sv; sv.kind = JVal::STR;
            if (!have_id) { sv.str = name->str + "-rule-" + std::to_string(i + 1); r.obj.push_back({"$ruleId", sv}); }
            if (!have_product) { sv.str = name->str; r.obj.push_back({"$productId", sv}); }
            defs.arr.push_back(r);
        }
        S rules_json; json_write(defs, &rules_json);
        return dx_codegen_impl(code, rules_json, query_xml, output_xml, name->str, res, err);
    } catch (std::exception& e) { *err = e.what(); return DX_ERR_PARSE; }
}

int dx_codegen_impl(const std::string& user_code, const std::string& rules_json, const std::string& query_xml, const std::string& output_xml,
                    const std::string& product, dx_codegen_result* res, std::string* err) {
    try {
        auto strip_bom = [](S s) { return s.compare(0, 3, "\xEF\xBB\xBF") == 0 ? s.substr(3) : s; };
        Engine E;
        E.code = strip_bom(user_code);
        E.rules = load_rules(strip_bom(rules_json));
        E.queries = load_templates(strip_bom(query_xml), "query", "type");
        E.outputs_t = load_templates(strip_bom(output_xml), "outputTemplate", "id");
        S code = E.run(product);
        if (getenv("DX_CODEGEN_DEBUG")) fprintf(stderr, "---- after the Process* passes ----\n%s\n----\n", code.c_str());
        // OUTPUT <table> TO <sinks>;   (Engine.cs:495-514)
        for (auto& m : matches(code, rx("OUTPUT\\s{1,}(.*?)\\s{1,}TO\\s{1,}([^;]*);"))) {
            for (auto& sink : split(m[2], ',')) res->outputs.push_back({m[1], trim(sink)});
            code = replace_all(code, m[0], "");
        }
        // CREATE TABLE <name> (<schema>);   (Engine.cs:558-577)
        for (auto& m : matches(code, rx("CREATE TABLE\\s{1,}(.*?)\\s{0,}\\((.*?)\\)\\s{0,};"))) {
            res->tables.push_back({m[1], m[2]});
            code = replace_all(code, m[0], "");
        }
        code = replace_all(code, "--DataXStates--", "");
        // ... FROM DataXProcessedInput TIMEWINDOW('<duration>') ...   (Engine.cs:594-626)
        for (auto& m : matches(code, rx("\\s{0,}--DataXQuery--\\s{0,}([^;]*?)FROM\\s{1,}([^\\s]+)(\\s{1,})TIMEWINDOW\\s{0,}\\(\\s{0,}(.*?)\\s{0,}\\)\\s{0,}([^;]*?)"))) {
            const S window = replace_all(trim(m[4]), "'", "");
            const S table = trim(m[2]) + "_" + replace_all(window, " ", "");
            if (lower(trim(m[2])) != "dataxprocessedinput") throw std::runtime_error("'DataXProcessedInput' is the only table for which the TIMEWINDOW can be specified");
            S q = std::regex_replace(m[0], std::regex("\\bDataXProcessedInput\\b", std::regex::icase), table);
            q = std::regex_replace(q, rx(trim(m[4]), false), "");
            q = std::regex_replace(q, rx("(TIMEWINDOW\\s{0,}\\(\\s{0,}\\)\\s{0,})"), "");
            bool have = false; for (auto& w : res->windows) if (w.first == table) have = true;
            if (!have) res->windows.push_back({table, window});
            code = replace_all(code, m[0], q);
        }
        for (auto& o : res->outputs) if (lower(trim(o.second)) == "metrics") res->metrics.push_back(o.first);
        // <query> WITH UPSERT <table>   (Engine.cs:580-592)
        for (auto& m : matches(code, rx("\\s{0,}--DataXQuery--\\s{0,}([^;]*)WITH\\s{1,}UPSERT\\s{1,}([^;]*)")))
            code = replace_all(code, m[0], "\r\n\n--DataXQuery--\r\n" + trim(m[2]) + " = " + trim(m[1]) + "\r\n");
        // pretty-print and tidy (Engine.cs:93-100, 533-555)
        S f = Printer(code).run();
        f = replace_all(f, "\n", "\r\n");
        f = replace_all(f, "--DataXQuery--", "\r\n--DataXQuery--\r\n");
        f = replace_all(f, "--Rule", "\r\n\r\n--Rule");
        f = replace_all(f, ";", "");
        f = trim(f); f = trim_chars(f, "\n\r\t");
        for (auto& m : matches(f, rx("--DataXQuery--\\s{0,}--DataXQuery--", false))) { f = replace_all(f, m[0], "--DataXQuery--"); f = trim_chars(f, "\n\r\t"); }
        if (ends_with(f, "--DataXQuery--")) f = f.substr(0, f.size() - 14);
        else if (ends_with(f, "--DataXQuery--\n")) f = f.substr(0, f.size() - 15) + "\n";
        res->code = trim(f);
        return DX_OK;
    } catch (std::exception& e) { *err = e.what(); return DX_ERR_PARSE; }
}
