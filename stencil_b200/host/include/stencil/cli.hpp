// cli.hpp — a small self-contained command-line parser.
//
// The reference parses its flags with CLI11 v2.4.1, fetched from GitHub at configure time
// (cmake/AddCLI11.cmake:11-21); there is no network here, so this header provides the
// handful of features program_options.cpp needs: "-s,--long" option names, required
// options, defaults, list-valued options, boolean flags, --help, and an error message +
// non-zero exit code on a parse error (reference behaviour: src/program_options.cpp:39-44,
// src/main.cpp:59-61).
#ifndef STENCIL_CLI_HPP
#define STENCIL_CLI_HPP

#include <cstdlib>
#include <functional>
#include <iostream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace cli {

struct ParseError : std::runtime_error {
    int exit_code;
    ParseError(std::string const& what, int code) : std::runtime_error(what), exit_code(code) {}
};

class Parser {
public:
    explicit Parser(std::string description = "") : description_(std::move(description)) {}

    struct Opt {
        std::vector<std::string> names;
        std::string help;
        bool is_flag = false, is_list = false, required = false, seen = false;
        std::string default_text;
        std::function<void(std::string const&)> assign;  // one value
        std::function<void()> reset_list;                // lists: clear before first value
        Opt& require() {
            required = true;
            return *this;
        }
    };

    template <class T>
    Opt& option(std::string const& names, T& target, std::string const& help) {
        Opt& o = add(names, help);
        o.assign = [&target, names](std::string const& v) { target = convert<T>(v, names); };
        return o;
    }
    template <class T>
    Opt& option(std::string const& names, T& target, std::string const& help, T default_value) {
        target = default_value;
        Opt& o = option(names, target, help);
        std::ostringstream os;
        os << default_value;
        o.default_text = os.str();
        return o;
    }
    template <class T>
    Opt& list(std::string const& names, std::vector<T>& target, std::string const& help) {
        Opt& o = add(names, help);
        o.is_list = true;
        o.reset_list = [&target] { target.clear(); };
        o.assign = [&target, names](std::string const& v) { target.push_back(convert<T>(v, names)); };
        return o;
    }
    Opt& flag(std::string const& names, bool& target, std::string const& help) {
        target = false;
        Opt& o = add(names, help);
        o.is_flag = true;
        o.assign = [&target](std::string const&) { target = true; };
        return o;
    }

    void parse(int argc, char** argv) {
        program_ = argc > 0 ? argv[0] : "program";
        Opt* open_list = nullptr;
        for (int i = 1; i < argc; ++i) {
            std::string arg = argv[i];
            if (arg == "-h" || arg == "--help") throw ParseError(usage(), 0);
            std::string inline_value;
            bool has_inline = false;
            if (arg.rfind("--", 0) == 0) {
                auto eq = arg.find('=');
                if (eq != std::string::npos) {
                    inline_value = arg.substr(eq + 1);
                    arg = arg.substr(0, eq);
                    has_inline = true;
                }
            }
            Opt* o = looks_like_option(arg) ? find(arg) : nullptr;
            if (!o) {
                if (looks_like_option(arg) && !is_number(arg))
                    throw ParseError("The following argument was not expected: " + arg, 109);
                if (open_list) {
                    open_list->assign(arg);
                    continue;
                }
                throw ParseError("The following argument was not expected: " + arg, 109);
            }
            open_list = nullptr;
            if (o->is_flag) {
                o->assign("");
                o->seen = true;
                continue;
            }
            if (o->is_list) {
                if (!o->seen) o->reset_list();
                o->seen = true;
                if (has_inline) o->assign(inline_value);
                open_list = o;
                continue;
            }
            std::string value;
            if (has_inline) {
                value = inline_value;
            } else {
                if (i + 1 >= argc) throw ParseError(o->names.front() + ": 1 required value missing", 113);
                value = argv[++i];
            }
            o->assign(value);
            o->seen = true;
        }
        for (auto const& o : opts_)
            if (o.required && !o.seen) throw ParseError(join(o.names) + " is required", 106);
    }

    // Mirrors CLI::App::exit: prints the message (help to stdout, errors to stderr) and
    // returns the process exit code.
    int exit(ParseError const& e) const {
        if (e.exit_code == 0) {
            std::cout << e.what();
        } else {
            std::cerr << e.what() << "\nRun with --help for more information.\n";
        }
        return e.exit_code;
    }

    std::string usage() const {
        std::ostringstream os;
        if (!description_.empty()) os << description_ << "\n";
        os << "Usage: " << program_ << " [OPTIONS]\n\nOptions:\n  -h,--help  Print this help message and exit\n";
        for (auto const& o : opts_) {
            os << "  " << join(o.names);
            if (!o.is_flag) os << (o.is_list ? " VALUE ..." : " VALUE");
            if (o.required) os << " REQUIRED";
            if (!o.default_text.empty()) os << " [" << o.default_text << "]";
            os << "  " << o.help << "\n";
        }
        return os.str();
    }

private:
    std::string description_, program_ = "program";
    std::vector<Opt> opts_;

    Opt& add(std::string const& names, std::string const& help) {
        Opt o;
        std::stringstream ss(names);
        std::string item;
        while (std::getline(ss, item, ',')) o.names.push_back(item);
        o.help = help;
        opts_.reserve(64);  // references handed out stay valid
        opts_.push_back(std::move(o));
        return opts_.back();
    }
    Opt* find(std::string const& name) {
        for (auto& o : opts_)
            for (auto const& n : o.names)
                if (n == name) return &o;
        return nullptr;
    }
    static bool looks_like_option(std::string const& s) { return s.size() >= 2 && s[0] == '-'; }
    static bool is_number(std::string const& s) {
        char* end = nullptr;
        std::strtod(s.c_str(), &end);
        return end && *end == '\0';
    }
    static std::string join(std::vector<std::string> const& v) {
        std::string out;
        for (auto const& s : v) out += (out.empty() ? "" : ",") + s;
        return out;
    }
    template <class T>
    static T convert(std::string const& v, std::string const& names) {
        if constexpr (std::is_same_v<T, std::string>) {
            return v;
        } else {
            std::istringstream is(v);
            T out{};
            if constexpr (std::is_unsigned_v<T>) {
                if (!v.empty() && v[0] == '-')
                    throw ParseError("Could not convert: " + names + " = " + v, 104);
            }
            is >> out;
            if (is.fail() || !is.eof()) throw ParseError("Could not convert: " + names + " = " + v, 104);
            return out;
        }
    }
};

}  // namespace cli
#endif  // STENCIL_CLI_HPP
