/* Minimal stand-in for <boost/program_options.hpp> (Boost is not installed in this image), just
 * enough for the reference's main() (DInDel.cpp:4069-4188) and getParameters (DInDel.cpp:3907-3975):
 * long options only (`--name value`, `--name=value`, valueless flags), typed defaults, count(),
 * operator[] + as<T>().  Test infrastructure only (oracle/_ref full-binary build); gnu++98. */
#ifndef DINDEL_B200_SHIM_BOOST_PROGRAM_OPTIONS_HPP
#define DINDEL_B200_SHIM_BOOST_PROGRAM_OPTIONS_HPP
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>
#include <stdint.h>

namespace boost { namespace program_options {

struct error { virtual ~error() {} };

struct value_base {
    bool has_default;
    std::string default_text;
    value_base() : has_default(false) {}
    virtual ~value_base() {}
};

template <class T> struct typed_value : value_base {
    typed_value *default_value(const T &v)
    {
        std::ostringstream os;
        os.precision(17);
        os << v;
        default_text = os.str();
        has_default = true;
        return this;
    }
};

template <class T> typed_value<T> *value() { return new typed_value<T>(); }

struct option_def {
    std::string name, help;
    value_base *sem; /* 0: flag without a value */
};

class options_description;

class options_adder {
    options_description *owner;
public:
    explicit options_adder(options_description *o) : owner(o) {}
    options_adder &operator()(const char *name, const char *help);
    options_adder &operator()(const char *name, value_base *sem, const char *help);
};

class options_description {
public:
    std::string caption;
    std::vector<option_def> opts;
    options_description() {}
    explicit options_description(const std::string &c) : caption(c) {}
    options_adder add_options() { return options_adder(this); }
    options_description &add(const options_description &o)
    {
        opts.insert(opts.end(), o.opts.begin(), o.opts.end());
        return *this;
    }
    const option_def *find(const std::string &n) const
    {
        for (size_t i = 0; i < opts.size(); i++)
            if (opts[i].name == n) return &opts[i];
        return 0;
    }
};

inline options_adder &options_adder::operator()(const char *name, const char *help)
{
    option_def d;
    d.name = name; d.help = help; d.sem = 0;
    owner->opts.push_back(d);
    return *this;
}
inline options_adder &options_adder::operator()(const char *name, value_base *sem, const char *help)
{
    option_def d;
    d.name = name; d.help = help; d.sem = sem;
    owner->opts.push_back(d);
    return *this;
}

inline std::ostream &operator<<(std::ostream &os, const options_description &d)
{
    os << d.caption << "\n";
    for (size_t i = 0; i < d.opts.size(); i++) {
        os << "  --" << d.opts[i].name;
        if (d.opts[i].sem) {
            os << " arg";
            if (d.opts[i].sem->has_default) os << " (=" << d.opts[i].sem->default_text << ")";
        }
        os << "\t" << d.opts[i].help << "\n";
    }
    return os;
}

struct variable_value {
    std::string text;
    template <class T> T as() const
    {
        std::istringstream is(text);
        T v = T();
        is >> v;
        if (is.fail()) throw error();
        return v;
    }
};
template <> inline std::string variable_value::as<std::string>() const { return text; }

struct parsed_options {
    std::vector<std::pair<std::string, std::string> > values;
    const options_description *desc;
};

inline parsed_options parse_command_line(int argc, char **argv, const options_description &desc)
{
    parsed_options p;
    p.desc = &desc;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        if (a.size() < 3 || a[0] != '-' || a[1] != '-') throw error();
        std::string name = a.substr(2), val;
        bool have_val = false;
        size_t eq = name.find('=');
        if (eq != std::string::npos) {
            val = name.substr(eq + 1);
            name = name.substr(0, eq);
            have_val = true;
        }
        const option_def *d = desc.find(name);
        if (!d) throw error();
        if (d->sem) {
            if (!have_val) {
                if (i + 1 >= argc) throw error();
                val = argv[++i];
            }
        } else if (have_val) throw error();
        p.values.push_back(std::make_pair(name, val));
    }
    return p;
}

class variables_map {
    std::map<std::string, variable_value> m;
public:
    size_t count(const std::string &n) const { return m.count(n); }
    const variable_value &operator[](const std::string &n) const
    {
        std::map<std::string, variable_value>::const_iterator it = m.find(n);
        if (it == m.end()) throw error();
        return it->second;
    }
    void set(const std::string &n, const std::string &v, bool overwrite)
    {
        if (overwrite || !m.count(n)) m[n].text = v;
    }
};

inline void store(const parsed_options &p, variables_map &vm)
{
    for (size_t i = 0; i < p.values.size(); i++) vm.set(p.values[i].first, p.values[i].second, true);
    for (size_t i = 0; i < p.desc->opts.size(); i++) {
        const option_def &d = p.desc->opts[i];
        if (d.sem && d.sem->has_default) vm.set(d.name, d.sem->default_text, false);
    }
}

inline void notify(variables_map &) {}

}} /* namespace boost::program_options */
#endif
