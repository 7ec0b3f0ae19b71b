// toml_subset.cpp — see toml_subset.hpp.
#include "toml_subset.hpp"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>

namespace tomls {

const Value* Table::find(const std::string& key) const {
    for (const auto& kv : items)
        if (kv.first == key) return &kv.second;
    return nullptr;
}
Value* Table::find(const std::string& key) {
    for (auto& kv : items)
        if (kv.first == key) return &kv.second;
    return nullptr;
}
Value& Table::insert(const std::string& key, Value v) {
    items.emplace_back(key, std::move(v));
    return items.back().second;
}

const char* Value::type_name() const {
    switch (kind) {
        case Kind::String: return "string";
        case Kind::Integer: return "integer";
        case Kind::Float: return "float";
        case Kind::Boolean: return "boolean";
        case Kind::Array: return "array";
        case Kind::Table: return "table";
    }
    return "?";
}

namespace {

struct Parser {
    const std::string& t;
    size_t p = 0;
    int line = 1;
    explicit Parser(const std::string& text) : t(text) {}

    [[noreturn]] void fail(const std::string& msg) const { throw ParseError(line, msg); }
    bool eof() const { return p >= t.size(); }
    char peek(size_t k = 0) const { return p + k < t.size() ? t[p + k] : '\0'; }
    bool starts(const char* s) const { return t.compare(p, std::strlen(s), s) == 0; }

    void skip_ws() {
        while (!eof() && (t[p] == ' ' || t[p] == '\t')) p++;
    }
    void skip_comment() {
        if (peek() == '#') {
            while (!eof() && t[p] != '\n') {
                unsigned char c = (unsigned char)t[p];
                if ((c < 0x20 && c != '\t' && c != '\r') || c == 0x7f) fail("control character in comment");
                p++;
            }
        }
    }
    bool eat_newline() {
        if (peek() == '\n') { p++; line++; return true; }
        if (peek() == '\r' && peek(1) == '\n') { p += 2; line++; return true; }
        return false;
    }
    // whitespace, comments and newlines (inside arrays, between statements)
    void skip_blank() {
        for (;;) {
            skip_ws();
            skip_comment();
            if (!eat_newline()) break;
        }
    }
    void end_of_statement() {
        skip_ws();
        skip_comment();
        if (eof()) return;
        if (!eat_newline()) fail(std::string("expected end of line, found '") + peek() + "'");
    }

    static bool bare_char(char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9') || c == '_' || c == '-'; }

    void append_utf8(std::string& out, uint32_t cp) {
        if (cp > 0x10ffff || (cp >= 0xd800 && cp <= 0xdfff)) fail("invalid unicode scalar in escape");
        if (cp < 0x80) out += (char)cp;
        else if (cp < 0x800) { out += (char)(0xc0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3f)); }
        else if (cp < 0x10000) { out += (char)(0xe0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3f)); out += (char)(0x80 | (cp & 0x3f)); }
        else { out += (char)(0xf0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3f)); out += (char)(0x80 | ((cp >> 6) & 0x3f)); out += (char)(0x80 | (cp & 0x3f)); }
    }

    void parse_escape(std::string& out, bool multiline) {
        // at the character after the backslash
        char c = peek();
        p++;
        switch (c) {
            case 'b': out += '\b'; return;
            case 't': out += '\t'; return;
            case 'n': out += '\n'; return;
            case 'f': out += '\f'; return;
            case 'r': out += '\r'; return;
            case '"': out += '"'; return;
            case '\\': out += '\\'; return;
            case 'u': case 'U': {
                int n = c == 'u' ? 4 : 8;
                uint32_t cp = 0;
                for (int k = 0; k < n; k++) {
                    char h = peek();
                    int d = (h >= '0' && h <= '9') ? h - '0' : (h >= 'a' && h <= 'f') ? h - 'a' + 10 : (h >= 'A' && h <= 'F') ? h - 'A' + 10 : -1;
                    if (d < 0) fail("bad \\u escape");
                    cp = cp * 16 + (uint32_t)d;
                    p++;
                }
                append_utf8(out, cp);
                return;
            }
            default:
                if (multiline && (c == ' ' || c == '\t' || c == '\n' || c == '\r')) {   // line-ending backslash
                    p--;
                    size_t save = p; int save_line = line;
                    skip_ws();
                    if (!eat_newline()) { p = save; line = save_line; fail("invalid escape"); }
                    while (!eof() && (peek() == ' ' || peek() == '\t' || peek() == '\n' || peek() == '\r')) {
                        if (!eat_newline()) p++;
                    }
                    return;
                }
                fail(std::string("invalid escape \\") + c);
        }
    }

    std::string parse_basic_string() {
        // at the opening quote
        std::string out;
        if (starts("\"\"\"")) {
            p += 3;
            eat_newline();
            for (;;) {
                if (eof()) fail("unterminated multi-line string");
                if (starts("\"\"\"")) {
                    p += 3;
                    // up to two extra quotes belong to the string
                    int extra = 0;
                    while (peek() == '"' && extra < 2) { out += '"'; p++; extra++; }
                    return out;
                }
                char c = peek();
                if (c == '\\') { p++; parse_escape(out, true); continue; }
                if (eat_newline()) { out += '\n'; continue; }
                if (((unsigned char)c < 0x20 && c != '\t') || c == 0x7f) fail("control character in string");
                out += c; p++;
            }
        }
        p++;
        for (;;) {
            if (eof() || peek() == '\n') fail("unterminated string");
            char c = peek();
            if (c == '"') { p++; return out; }
            if (c == '\\') { p++; parse_escape(out, false); continue; }
            if (((unsigned char)c < 0x20 && c != '\t') || c == 0x7f) fail("control character in string");
            out += c; p++;
        }
    }

    std::string parse_literal_string() {
        std::string out;
        if (starts("'''")) {
            p += 3;
            eat_newline();
            for (;;) {
                if (eof()) fail("unterminated multi-line literal string");
                if (starts("'''")) {
                    p += 3;
                    int extra = 0;
                    while (peek() == '\'' && extra < 2) { out += '\''; p++; extra++; }
                    return out;
                }
                if (eat_newline()) { out += '\n'; continue; }
                char c = peek();
                if (((unsigned char)c < 0x20 && c != '\t') || c == 0x7f) fail("control character in string");
                out += c; p++;
            }
        }
        p++;
        for (;;) {
            if (eof() || peek() == '\n') fail("unterminated literal string");
            char c = peek();
            if (c == '\'') { p++; return out; }
            if (((unsigned char)c < 0x20 && c != '\t') || c == 0x7f) fail("control character in string");
            out += c; p++;
        }
    }

    std::string parse_key_part() {
        skip_ws();
        if (peek() == '"') {
            if (starts("\"\"\"")) fail("multi-line string as key");
            return parse_basic_string();
        }
        if (peek() == '\'') {
            if (starts("'''")) fail("multi-line string as key");
            return parse_literal_string();
        }
        size_t b = p;
        while (!eof() && bare_char(peek())) p++;
        if (p == b) fail(std::string("expected a key, found '") + peek() + "'");
        return t.substr(b, p - b);
    }

    std::vector<std::string> parse_key() {
        std::vector<std::string> parts;
        parts.push_back(parse_key_part());
        for (;;) {
            skip_ws();
            if (peek() != '.') break;
            p++;
            parts.push_back(parse_key_part());
        }
        return parts;
    }

    Value parse_value();
    Value parse_number_or_date();
    Value parse_array();
    Value parse_inline_table();
    void assign(Table& tab, const std::vector<std::string>& key, Value v);
    Table parse_document();
};

Value Parser::parse_number_or_date() {
    size_t b = p;
    while (!eof()) {
        char c = peek();
        if ((c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c == '+' || c == '-' || c == '.' || c == '_' || c == ':') p++;
        else break;
    }
    std::string tok = t.substr(b, p - b);
    if (tok.empty()) fail(std::string("expected a value, found '") + peek() + "'");
    if (tok == "true") return Value::boolean(true);
    if (tok == "false") return Value::boolean(false);
    std::string body = tok;
    int sign = 1;
    if (body[0] == '+' || body[0] == '-') { sign = body[0] == '-' ? -1 : 1; body = body.substr(1); }
    if (body == "inf") return Value::floating(sign * std::numeric_limits<double>::infinity());
    if (body == "nan") return Value::floating(std::numeric_limits<double>::quiet_NaN());
    if (body.empty()) fail("bad number '" + tok + "'");
    // dates / times: digits followed by '-' digits '-' or ':' patterns
    if (body.size() >= 5 && ((body[4] == '-' && std::isdigit((unsigned char)body[0])) || (body.size() > 2 && body[2] == ':')))
        fail("dates and times are not supported by this TOML reader ('" + tok + "')");
    auto strip_underscores = [&](const std::string& s, auto is_digit) {
        std::string o;
        for (size_t k = 0; k < s.size(); k++) {
            if (s[k] == '_') {
                if (k == 0 || k + 1 >= s.size() || !is_digit(s[k - 1]) || !is_digit(s[k + 1])) fail("misplaced underscore in number '" + tok + "'");
                continue;
            }
            o += s[k];
        }
        return o;
    };
    auto dec = [](char c) { return c >= '0' && c <= '9'; };
    if (body.size() > 2 && body[0] == '0' && (body[1] == 'x' || body[1] == 'o' || body[1] == 'b')) {
        if (tok[0] == '+' || tok[0] == '-') fail("sign on a prefixed integer '" + tok + "'");
        int base = body[1] == 'x' ? 16 : body[1] == 'o' ? 8 : 2;
        auto ok = [&](char c) {
            if (base == 16) return dec(c) || (c >= 'a' && c <= 'f') || (c >= 'A' && c <= 'F');
            if (base == 8) return c >= '0' && c <= '7';
            return c == '0' || c == '1';
        };
        std::string digits = strip_underscores(body.substr(2), ok);
        if (digits.empty()) fail("bad integer '" + tok + "'");
        for (char c : digits) if (!ok(c)) fail("bad digit in integer '" + tok + "'");
        errno = 0;
        unsigned long long v = std::strtoull(digits.c_str(), nullptr, base);
        if (errno || v > (unsigned long long)std::numeric_limits<int64_t>::max()) fail("integer out of range '" + tok + "'");
        return Value::integer((int64_t)v);
    }
    bool is_float = body.find('.') != std::string::npos || body.find('e') != std::string::npos || body.find('E') != std::string::npos;
    std::string clean = strip_underscores(body, dec);
    // validate shape: digits [ . digits ] [ e [+-] digits ]
    size_t k = 0;
    auto digits_run = [&]() { size_t s0 = k; while (k < clean.size() && dec(clean[k])) k++; return k - s0; };
    size_t int_digits = digits_run();
    if (int_digits == 0) fail("bad number '" + tok + "'");
    if (int_digits > 1 && clean[0] == '0') fail("leading zero in number '" + tok + "'");
    if (k < clean.size() && clean[k] == '.') { k++; if (digits_run() == 0) fail("bad float '" + tok + "'"); }
    if (k < clean.size() && (clean[k] == 'e' || clean[k] == 'E')) {
        k++;
        if (k < clean.size() && (clean[k] == '+' || clean[k] == '-')) k++;
        if (digits_run() == 0) fail("bad exponent in '" + tok + "'");
    }
    if (k != clean.size()) fail("bad number '" + tok + "'");
    if (is_float) return Value::floating(sign * std::strtod(clean.c_str(), nullptr));
    errno = 0;
    unsigned long long mag = std::strtoull(clean.c_str(), nullptr, 10);
    if (errno || mag > (unsigned long long)std::numeric_limits<int64_t>::max() + (sign < 0 ? 1ull : 0ull)) fail("integer out of range '" + tok + "'");
    return Value::integer(sign < 0 ? (int64_t)(0ull - mag) : (int64_t)mag);
}

Value Parser::parse_array() {
    p++;   // [
    Value arr = Value::array();
    for (;;) {
        skip_blank();
        if (eof()) fail("unterminated array");
        if (peek() == ']') { p++; break; }
        arr.arr->push_back(parse_value());
        skip_blank();
        if (peek() == ',') { p++; continue; }
        if (peek() == ']') { p++; break; }
        fail("expected ',' or ']' in array");
    }
    arr.inline_closed = true;
    return arr;
}

Value Parser::parse_inline_table() {
    p++;   // {
    Value tab = Value::table();
    skip_ws();
    if (peek() == '}') { p++; tab.inline_closed = true; return tab; }
    for (;;) {
        skip_ws();
        std::vector<std::string> key = parse_key();
        skip_ws();
        if (peek() != '=') fail("expected '=' in inline table");
        p++;
        skip_ws();
        assign(*tab.tab, key, parse_value());
        skip_ws();
        if (peek() == ',') { p++; continue; }
        if (peek() == '}') { p++; break; }
        fail("expected ',' or '}' in inline table");
    }
    // close the table and everything defined inside it
    tab.inline_closed = true;
    return tab;
}

Value Parser::parse_value() {
    skip_ws();
    char c = peek();
    if (c == '"') return Value::string(parse_basic_string());
    if (c == '\'') return Value::string(parse_literal_string());
    if (c == '[') return parse_array();
    if (c == '{') return parse_inline_table();
    return parse_number_or_date();
}

// key = value inside `tab`; dotted keys create intermediate tables.
void Parser::assign(Table& tab, const std::vector<std::string>& key, Value v) {
    Table* cur = &tab;
    for (size_t k = 0; k + 1 < key.size(); k++) {
        Value* nxt = cur->find(key[k]);
        if (!nxt) nxt = &cur->insert(key[k], Value::table());
        else if (!nxt->is_table() || nxt->inline_closed || nxt->array_of_tables) fail("key '" + key[k] + "' is not an open table");
        cur = nxt->tab.get();
    }
    if (cur->contains(key.back())) fail("duplicate key '" + key.back() + "'");
    cur->insert(key.back(), std::move(v));
}

Table Parser::parse_document() {
    Table root;
    Table* cur = &root;
    for (;;) {
        skip_blank();
        if (eof()) break;
        if (peek() == '[') {
            bool aot = peek(1) == '[';
            p += aot ? 2 : 1;
            std::vector<std::string> key = parse_key();
            skip_ws();
            if (aot) { if (!starts("]]")) fail("expected ']]'"); p += 2; }
            else { if (peek() != ']') fail("expected ']'"); p++; }
            end_of_statement();
            Table* at = &root;
            for (size_t k = 0; k + 1 < key.size(); k++) {
                Value* nxt = at->find(key[k]);
                if (!nxt) nxt = &at->insert(key[k], Value::table());
                if (nxt->is_array() && nxt->array_of_tables) at = nxt->arr->back().tab.get();
                else if (nxt->is_table() && !nxt->inline_closed) at = nxt->tab.get();
                else fail("'" + key[k] + "' is not a table");
            }
            const std::string& last = key.back();
            Value* ex = at->find(last);
            if (aot) {
                if (!ex) { Value a = Value::array(); a.array_of_tables = true; ex = &at->insert(last, std::move(a)); }
                else if (!ex->is_array() || !ex->array_of_tables) fail("'" + last + "' is not an array of tables");
                ex->arr->push_back(Value::table());
                cur = ex->arr->back().tab.get();
            } else {
                if (!ex) { Value tb = Value::table(); tb.defined_by_header = true; ex = &at->insert(last, std::move(tb)); }
                else {
                    if (!ex->is_table() || ex->inline_closed || ex->defined_by_header) fail("table '" + last + "' defined twice");
                    ex->defined_by_header = true;
                }
                cur = ex->tab.get();
            }
            continue;
        }
        std::vector<std::string> key = parse_key();
        skip_ws();
        if (peek() != '=') fail("expected '=' after key");
        p++;
        Value v = parse_value();
        assign(*cur, key, std::move(v));
        end_of_statement();
    }
    return root;
}

void json_string(std::string& out, const std::string& s) {
    out += '"';
    for (unsigned char c : s) {
        if (c == '"') out += "\\\"";
        else if (c == '\\') out += "\\\\";
        else if (c == '\n') out += "\\n";
        else if (c == '\r') out += "\\r";
        else if (c == '\t') out += "\\t";
        else if (c < 0x20 || c == 0x7f) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); out += b; }
        else out += (char)c;
    }
    out += '"';
}

void json_value(std::string& out, const Value& v) {
    switch (v.kind) {
        case Kind::String: json_string(out, v.s); break;
        case Kind::Integer: out += std::to_string(v.i); break;
        case Kind::Float: {
            if (std::isnan(v.f)) out += "NaN";
            else if (std::isinf(v.f)) out += v.f > 0 ? "Infinity" : "-Infinity";
            else { char b[40]; std::snprintf(b, sizeof b, "%.17g", v.f); out += b; if (!std::strpbrk(b, ".eE")) out += ".0"; }
            break;
        }
        case Kind::Boolean: out += v.b ? "true" : "false"; break;
        case Kind::Array: {
            out += '[';
            for (size_t k = 0; k < v.arr->size(); k++) { if (k) out += ", "; json_value(out, (*v.arr)[k]); }
            out += ']';
            break;
        }
        case Kind::Table: {
            out += '{';
            bool first = true;
            for (const auto& kv : v.tab->items) {
                if (!first) out += ", ";
                first = false;
                json_string(out, kv.first);
                out += ": ";
                json_value(out, kv.second);
            }
            out += '}';
            break;
        }
    }
}

}  // namespace

Table parse(const std::string& text) {
    Parser ps(text);
    return ps.parse_document();
}

std::string to_json(const Value& v) {
    std::string out;
    json_value(out, v);
    return out;
}

std::string to_json(const Table& t) {
    Value v = Value::table();
    *v.tab = t;
    return to_json(v);
}

}  // namespace tomls
