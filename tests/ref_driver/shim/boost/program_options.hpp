// TEST INFRASTRUCTURE.  A minimal stand-in for <boost/program_options.hpp> (boost is not installed in this image): just
// the subset examples/shared/ArgParser.h uses, so that the reference's UNMODIFIED example drivers compile against
// include/Opt.h + libOptB200.so (tests/ref_driver/Makefile).  Options: --name value, --name=value, --flag (implicit
// value), and `name = value` lines of a config file; earlier stores win, as in boost.
#pragma once
#include <cstdlib>
#include <iostream>
#include <istream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace boost { namespace program_options {

struct value_base {
    bool has_default = false, has_implicit = false;
    std::string def, implicit;
    virtual ~value_base() {}
};
template <class T> std::string to_text(const T& v) { std::ostringstream o; o << v; return o.str(); }
template <class T> struct typed_value : value_base {
    typed_value* default_value(const T& v) { has_default = true; def = to_text(v); return this; }
    typed_value* implicit_value(const T& v) { has_implicit = true; implicit = to_text(v); return this; }
};
template <class T> typed_value<T>* value() { return new typed_value<T>(); }

struct option_description { std::string name, help; std::shared_ptr<value_base> sem; };

class options_description;
class options_easy_init {
public:
    explicit options_easy_init(options_description* o) : owner(o) {}
    options_easy_init& operator()(const char* name, const char* help);
    options_easy_init& operator()(const char* name, value_base* v, const char* help);
private:
    options_description* owner;
};
class options_description {
public:
    explicit options_description(const std::string& caption = "") : caption_(caption) {}
    options_easy_init add_options() { return options_easy_init(this); }
    const option_description* find(const std::string& n) const { for (auto& o : opts) if (o.name == n) return &o; return nullptr; }
    std::vector<option_description> opts;
    std::string caption_;
};
inline options_easy_init& options_easy_init::operator()(const char* name, const char* help) {
    owner->opts.push_back({name, help, nullptr}); return *this;
}
inline options_easy_init& options_easy_init::operator()(const char* name, value_base* v, const char* help) {
    owner->opts.push_back({name, help, std::shared_ptr<value_base>(v)}); return *this;
}
inline std::ostream& operator<<(std::ostream& os, const options_description& d) {
    os << d.caption_ << ":\n";
    for (auto& o : d.opts) os << "  --" << o.name << "  " << o.help << "\n";
    return os;
}

struct parsed_options { std::vector<std::pair<std::string, std::string>> kv; const options_description* desc; };

class variable_value {
public:
    variable_value() {}
    explicit variable_value(const std::string& t) : text(t) {}
    template <class T> T as() const { T v{}; std::istringstream i(text); i >> std::boolalpha; conv(i, v); return v; }
private:
    template <class T> static void conv(std::istringstream& i, T& v) { i >> v; }
    static void conv(std::istringstream& i, std::string& v) { v = i.str(); }
    static void conv(std::istringstream& i, bool& v) {
        const std::string s = i.str();
        v = (s == "1" || s == "true" || s == "True" || s == "on" || s == "yes");
    }
    std::string text;
};
class variables_map : public std::map<std::string, variable_value> {
public:
    // an option that no store() set explicitly reads its default (kept under a shadow key so that a later store() -- the
    // config file after the command line -- can still override it, as in boost)
    const variable_value& operator[](const std::string& k) const {
        auto it = find(k);
        if (it == end()) it = find("\x01" + k);
        if (it == end()) throw std::runtime_error("program_options shim: no option '" + k + "'");
        return it->second;
    }
    const options_description* desc = nullptr;
};

inline parsed_options parse_command_line(int argc, const char* const argv[], const options_description& d) {
    parsed_options p; p.desc = &d;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a.compare(0, 2, "--") != 0) continue;
        a = a.substr(2);
        std::string key = a, val; bool has = false;
        const size_t eq = a.find('=');
        if (eq != std::string::npos) { key = a.substr(0, eq); val = a.substr(eq + 1); has = true; }
        const option_description* o = d.find(key);
        if (!o) throw std::runtime_error("program_options shim: unknown option --" + key);
        if (!has) {
            const bool next_is_value = i + 1 < argc && std::string(argv[i + 1]).compare(0, 2, "--") != 0;
            if (o->sem && o->sem->has_implicit && !next_is_value) val = o->sem->implicit;
            else if (o->sem && next_is_value) val = argv[++i];
            else val = "1";
        }
        p.kv.push_back({key, val});
    }
    return p;
}
inline parsed_options parse_config_file(std::istream& is, const options_description& d) {
    parsed_options p; p.desc = &d;
    std::string line;
    auto trim = [](std::string s) { const size_t a = s.find_first_not_of(" \t\r"), b = s.find_last_not_of(" \t\r"); return a == std::string::npos ? std::string() : s.substr(a, b - a + 1); };
    while (std::getline(is, line)) {
        const size_t h = line.find('#'); if (h != std::string::npos) line = line.substr(0, h);
        const size_t eq = line.find('=');
        if (eq == std::string::npos) continue;
        const std::string k = trim(line.substr(0, eq)), v = trim(line.substr(eq + 1));
        if (!k.empty() && d.find(k)) p.kv.push_back({k, v});
    }
    return p;
}
inline void store(const parsed_options& p, variables_map& m) {
    m.desc = p.desc;
    for (auto& kv : p.kv) if (!m.count(kv.first)) m.insert({kv.first, variable_value(kv.second)});     // first store wins
    for (auto& o : p.desc->opts)
        if (o.sem && o.sem->has_default && !m.count("\x01" + o.name)) m.insert({"\x01" + o.name, variable_value(o.sem->def)});
}
inline void notify(variables_map&) {}
}}  // namespace boost::program_options

