// Minimal stand-in for <boost/program_options.hpp>, TEST INFRASTRUCTURE ONLY.
//
// The reference's input.cxx is compiled UNCHANGED (in place, from /root/reference)
// to build the oracle; this image has no Boost headers, so this header supplies the
// small subset of Boost.ProgramOptions that input.cxx uses (input.cxx:8-9,16-929,
// 910-929,971-979,1503-1519): options_description / add_options / value<T>(&x) with
// default_value() and required(), variables_map with count() and operator[].as<T>(),
// parse_config_file (INI: [section] prefixes, '#' comments, key = value), store,
// notify and the multiple_occurrences exception.  It is only ever put on the include
// path by oracle/Makefile when a real Boost is not found.
#ifndef DES_ORACLE_COMPAT_BOOST_PROGRAM_OPTIONS_HPP
#define DES_ORACLE_COMPAT_BOOST_PROGRAM_OPTIONS_HPP

#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace boost { namespace program_options {

class error : public std::logic_error {
public:
    explicit error(const std::string &w) : std::logic_error(w) {}
};

class multiple_occurrences : public error {
    std::string name_;
public:
    explicit multiple_occurrences(const std::string &n)
        : error("option cannot be specified more than once"), name_(n) {}
    ~multiple_occurrences() throw() {}
    std::string get_option_name() const { return name_; }
};

namespace detail {
    inline std::string trim(const std::string &s) {
        const char *ws = " \t\r\n";
        std::string::size_type b = s.find_first_not_of(ws);
        if (b == std::string::npos) return std::string();
        std::string::size_type e = s.find_last_not_of(ws);
        return s.substr(b, e - b + 1);
    }
    template <class T> inline void convert(const std::string &name, const std::string &txt, T &out) {
        std::istringstream is(txt);
        is >> out;
        if (is.fail() || !(is >> std::ws).eof())
            throw error("the argument ('" + txt + "') for option '" + name + "' is invalid");
    }
    template <> inline void convert<std::string>(const std::string &, const std::string &txt, std::string &out) {
        out = txt;
    }
    template <> inline void convert<double>(const std::string &name, const std::string &txt, double &out) {
        char *end = 0;
        out = std::strtod(txt.c_str(), &end);
        if (txt.empty() || *end != '\0')
            throw error("the argument ('" + txt + "') for option '" + name + "' is invalid");
    }
    template <> inline void convert<bool>(const std::string &name, const std::string &txt, bool &out) {
        std::string s;
        for (std::string::size_type i = 0; i < txt.size(); ++i) s += char(std::tolower(txt[i]));
        if (s.empty() || s == "on" || s == "yes" || s == "1" || s == "true") out = true;
        else if (s == "off" || s == "no" || s == "0" || s == "false") out = false;
        else throw error("the argument ('" + txt + "') for option '" + name + "' is invalid");
    }
}

// Type-erased holder for one option's value.
class variable_value {
public:
    struct holder_base { virtual ~holder_base() {} };
    template <class T> struct holder : holder_base { T v; explicit holder(const T &x) : v(x) {} };
    std::shared_ptr<holder_base> h;
    bool defaulted;
    variable_value() : defaulted(false) {}
    template <class T> const T &as() const {
        const holder<T> *p = dynamic_cast<const holder<T>*>(h.get());
        if (!p) throw error("bad any_cast in variable_value::as");
        return p->v;
    }
    bool empty() const { return !h; }
};

class value_semantic {
public:
    virtual ~value_semantic() {}
    virtual void parse(const std::string &name, const std::string &txt, variable_value &vv) const = 0;
    virtual bool apply_default(variable_value &vv) const = 0;
    virtual void notify(const variable_value &vv) const = 0;
    virtual bool is_required() const = 0;
};

template <class T>
class typed_value : public value_semantic {
    T *store_;
    bool has_default_;
    T default_;
    bool required_;
public:
    explicit typed_value(T *p) : store_(p), has_default_(false), default_(), required_(false) {}
    template <class U> typed_value *default_value(const U &v) { default_ = T(v); has_default_ = true; return this; }
    typed_value *required() { required_ = true; return this; }
    void parse(const std::string &name, const std::string &txt, variable_value &vv) const {
        T v;
        detail::convert<T>(name, txt, v);
        vv.h.reset(new variable_value::holder<T>(v));
        vv.defaulted = false;
    }
    bool apply_default(variable_value &vv) const {
        if (!has_default_) return false;
        vv.h.reset(new variable_value::holder<T>(default_));
        vv.defaulted = true;
        return true;
    }
    void notify(const variable_value &vv) const { if (store_) *store_ = vv.as<T>(); }
    bool is_required() const { return required_; }
};

template <class T> inline typed_value<T> *value() { return new typed_value<T>(0); }
template <class T> inline typed_value<T> *value(T *p) { return new typed_value<T>(p); }

struct option_description {
    std::string name, help;
    std::shared_ptr<const value_semantic> sem;
};

class options_description;
class options_description_easy_init {
    options_description *owner_;
public:
    explicit options_description_easy_init(options_description *o) : owner_(o) {}
    options_description_easy_init &operator()(const char *name, const value_semantic *s, const char *help = "");
};

class options_description {
public:
    std::string caption;
    std::vector<option_description> opts;
    explicit options_description(const std::string &c = std::string()) : caption(c) {}
    options_description_easy_init add_options() { return options_description_easy_init(this); }
    const option_description *find(const std::string &n) const {
        for (std::size_t i = 0; i < opts.size(); ++i) if (opts[i].name == n) return &opts[i];
        return 0;
    }
};

inline options_description_easy_init &
options_description_easy_init::operator()(const char *name, const value_semantic *s, const char *help) {
    option_description d;
    d.name = name; d.help = help ? help : ""; d.sem.reset(s);
    owner_->opts.push_back(d);
    return *this;
}

inline std::ostream &operator<<(std::ostream &os, const options_description &d) {
    os << d.caption << ":\n";
    for (std::size_t i = 0; i < d.opts.size(); ++i)
        os << "  --" << d.opts[i].name << " arg\n      " << d.opts[i].help << "\n";
    return os;
}

struct parsed_options {
    const options_description *desc;
    std::vector<std::pair<std::string, std::string> > kv;
};

template <class charT>
inline parsed_options parse_config_file(const char *filename, const options_description &desc,
                                        bool allow_unregistered = false) {
    std::ifstream in(filename);
    if (!in) throw error(std::string("can not read options configuration file '") + filename + "'");
    parsed_options out;
    out.desc = &desc;
    std::string line, prefix;
    while (std::getline(in, line)) {
        std::string::size_type hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        line = detail::trim(line);
        if (line.empty()) continue;
        if (line[0] == '[' && line[line.size() - 1] == ']') {
            prefix = line.substr(1, line.size() - 2);
            if (!prefix.empty() && prefix[prefix.size() - 1] != '.') prefix += '.';
            continue;
        }
        std::string::size_type eq = line.find('=');
        if (eq == std::string::npos) throw error("the options configuration file contains an invalid line '" + line + "'");
        std::string name = prefix + detail::trim(line.substr(0, eq));
        std::string val = detail::trim(line.substr(eq + 1));
        if (!desc.find(name)) {
            if (allow_unregistered) continue;
            throw error("unrecognised option '" + name + "'");
        }
        out.kv.push_back(std::make_pair(name, val));
    }
    return out;
}

class variables_map : public std::map<std::string, variable_value> {
public:
    const options_description *desc;
    variables_map() : desc(0) {}
    const variable_value &operator[](const std::string &n) const {
        static const variable_value empty;
        const_iterator it = find(n);
        return it == end() ? empty : it->second;
    }
    std::size_t count(const std::string &n) const {
        return std::map<std::string, variable_value>::count(n);
    }
};

inline void store(const parsed_options &p, variables_map &vm) {
    vm.desc = p.desc;
    std::map<std::string, variable_value> &m = vm;
    std::map<std::string, bool> seen;
    for (std::size_t i = 0; i < p.kv.size(); ++i) {
        const std::string &name = p.kv[i].first;
        if (seen[name]) throw multiple_occurrences(name);
        seen[name] = true;
        const option_description *d = p.desc->find(name);
        std::map<std::string, variable_value>::iterator it = m.find(name);
        if (it != m.end() && !it->second.defaulted) continue;  // first store wins
        d->sem->parse(name, p.kv[i].second, m[name]);
    }
    for (std::size_t i = 0; i < p.desc->opts.size(); ++i) {
        const option_description &d = p.desc->opts[i];
        if (m.find(d.name) != m.end()) continue;
        variable_value vv;
        if (d.sem->apply_default(vv)) m[d.name] = vv;
    }
}

inline void notify(variables_map &vm) {
    if (!vm.desc) return;
    const std::map<std::string, variable_value> &m = vm;
    for (std::size_t i = 0; i < vm.desc->opts.size(); ++i) {
        const option_description &d = vm.desc->opts[i];
        std::map<std::string, variable_value>::const_iterator it = m.find(d.name);
        if (it == m.end()) {
            if (d.sem->is_required())
                throw error("the option '" + d.name + "' is required but missing");
            continue;
        }
        d.sem->notify(it->second);
    }
}

}} // namespace boost::program_options

#endif
