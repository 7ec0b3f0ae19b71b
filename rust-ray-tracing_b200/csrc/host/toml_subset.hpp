// toml_subset.hpp — a small TOML reader for the native host (the reference parses scene.toml with the `toml` crate,
// src/main.rs:36-39).  Supports what scene files use and a good deal more: comments, bare / quoted / dotted keys,
// [tables], [[arrays of tables]], inline tables, (nested, multi-line) arrays, basic and literal strings (single and
// multi-line), integers (decimal, hex, octal, binary, underscores), floats (fraction, exponent, inf, nan), booleans.
// Not supported: dates and times (reported as an error).  Errors carry line numbers.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace tomls {

struct ParseError : std::runtime_error {
    int line;
    ParseError(int line_, const std::string& msg) : std::runtime_error("TOML parse error at line " + std::to_string(line_) + ": " + msg), line(line_) {}
};

struct Value;
using Array = std::vector<Value>;
// Key order is kept (the reference's `toml::Table` iterates in sorted order; materials are looked up by name, so
// order only matters for diagnostics) — a vector of pairs plus lookup helpers.
struct Table {
    std::vector<std::pair<std::string, Value>> items;
    const Value* find(const std::string& key) const;
    Value* find(const std::string& key);
    Value& insert(const std::string& key, Value v);
    bool contains(const std::string& key) const { return find(key) != nullptr; }
};

enum class Kind { String, Integer, Float, Boolean, Array, Table };

struct Value {
    Kind kind = Kind::Integer;
    std::string s;
    int64_t i = 0;
    double f = 0.0;
    bool b = false;
    std::shared_ptr<Array> arr;
    std::shared_ptr<Table> tab;
    // bookkeeping for the TOML rules on redefinition
    bool defined_by_header = false;   // table created by a [header]
    bool inline_closed = false;       // inline table / static array: may not be extended
    bool array_of_tables = false;

    static Value string(std::string v) { Value x; x.kind = Kind::String; x.s = std::move(v); return x; }
    static Value integer(int64_t v) { Value x; x.kind = Kind::Integer; x.i = v; return x; }
    static Value floating(double v) { Value x; x.kind = Kind::Float; x.f = v; return x; }
    static Value boolean(bool v) { Value x; x.kind = Kind::Boolean; x.b = v; return x; }
    static Value array() { Value x; x.kind = Kind::Array; x.arr = std::make_shared<Array>(); return x; }
    static Value table() { Value x; x.kind = Kind::Table; x.tab = std::make_shared<Table>(); return x; }

    bool is_string() const { return kind == Kind::String; }
    bool is_integer() const { return kind == Kind::Integer; }
    bool is_float() const { return kind == Kind::Float; }
    bool is_boolean() const { return kind == Kind::Boolean; }
    bool is_array() const { return kind == Kind::Array; }
    bool is_table() const { return kind == Kind::Table; }
    const char* type_name() const;
};

// Parses a whole document; throws ParseError.
Table parse(const std::string& text);

// Canonical JSON rendering (floats with 17 significant digits) — used to cross-check against Python's tomllib.
std::string to_json(const Value& v);
std::string to_json(const Table& t);

}  // namespace tomls
