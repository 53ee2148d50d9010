// TEST SCAFFOLDING ONLY (oracle build). Header-only stand-in for the subset of
// boost::program_options that the reference's configuration layer uses
// (core/src/include/Configured.hpp:73-124, core/src/Configurator.cpp:18-60,
// core/src/ConfiguredModule.cpp:19-56, core/src/include/EnumWrapper.hpp:101-112).
// Boost is not installed in this image. This file exists only so that the reference's
// own thermodynamics sources compile unmodified into oracle/_ref; it is never linked
// into the product library.
//
// Semantics kept from real Boost because the reference relies on them:
//  * store(): the first *explicit* value stored for a key wins; a later explicit value
//    does not replace it, but an explicit value does replace one that was only defaulted.
//  * parse_config_file(): INI syntax, "[section]" + "key = value" => "section.key",
//    '#' comments, unknown keys ignored when allow_unregistered is true.
#pragma once
#include <any>
#include <istream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace boost {
namespace program_options {

    class validation_error : public std::logic_error {
    public:
        enum kind_t { invalid_option = 30 };
        explicit validation_error(kind_t)
            : std::logic_error("validation error")
        {
        }
    };

    class value_semantic {
    public:
        virtual ~value_semantic() = default;
        virtual bool has_default() const = 0;
        virtual std::any default_any() const = 0;
        virtual std::any parse(const std::string& text) const = 0;
    };

    namespace detail {
        template <typename T> inline T convert(const std::string& text)
        {
            std::istringstream is(text);
            T v {};
            is >> v;
            if (is.fail())
                throw validation_error(validation_error::invalid_option);
            return v;
        }
        template <> inline std::string convert<std::string>(const std::string& text)
        {
            return text;
        }
        template <> inline bool convert<bool>(const std::string& text)
        {
            std::string s;
            for (char c : text)
                s += (char)std::tolower((unsigned char)c);
            if (s == "1" || s == "true" || s == "yes" || s == "on")
                return true;
            if (s == "0" || s == "false" || s == "no" || s == "off")
                return false;
            throw validation_error(validation_error::invalid_option);
        }
    }

    template <typename T> class typed_value : public value_semantic {
    public:
        typed_value* default_value(const T& v)
        {
            m_default = v;
            m_hasDefault = true;
            return this;
        }
        typed_value* multitoken() { return this; }
        bool has_default() const override { return m_hasDefault; }
        std::any default_any() const override { return std::any(m_default); }
        std::any parse(const std::string& text) const override
        {
            return std::any(detail::convert<T>(text));
        }

    private:
        T m_default {};
        bool m_hasDefault = false;
    };

    template <typename T> inline typed_value<T>* value() { return new typed_value<T>(); }

    struct option_description {
        std::string name;
        std::shared_ptr<const value_semantic> semantic;
    };

    class options_description;
    class options_description_easy_init {
    public:
        explicit options_description_easy_init(options_description* o)
            : owner(o)
        {
        }
        options_description_easy_init& operator()(const char* name, const char* help);
        options_description_easy_init& operator()(
            const char* name, const value_semantic* s, const char* help = "");

    private:
        options_description* owner;
    };

    class options_description {
    public:
        options_description() = default;
        explicit options_description(const std::string&) { }
        options_description_easy_init add_options()
        {
            return options_description_easy_init(this);
        }
        const option_description* find(const std::string& name) const
        {
            for (const auto& o : m_options)
                if (o.name == name)
                    return &o;
            return nullptr;
        }
        const std::vector<option_description>& all() const { return m_options; }
        std::vector<option_description> m_options;
    };

    inline options_description_easy_init& options_description_easy_init::operator()(
        const char* name, const char*)
    {
        owner->m_options.push_back({ name, nullptr });
        return *this;
    }
    inline options_description_easy_init& options_description_easy_init::operator()(
        const char* name, const value_semantic* s, const char*)
    {
        owner->m_options.push_back({ name, std::shared_ptr<const value_semantic>(s) });
        return *this;
    }

    struct basic_option {
        std::string string_key;
        std::vector<std::string> value;
        bool unregistered = false;
    };

    struct parsed_options {
        const options_description* description = nullptr;
        std::vector<basic_option> options;
    };

    namespace command_line_style {
        enum style_t { unix_style = 1, allow_long_disguise = 2 };
    }

    class command_line_parser {
    public:
        command_line_parser(int argc, char** argv)
        {
            for (int k = 1; k < argc; ++k)
                if (argv && argv[k])
                    m_args.emplace_back(argv[k]);
        }
        command_line_parser& options(const options_description& d)
        {
            m_desc = &d;
            return *this;
        }
        command_line_parser& style(int) { return *this; }
        command_line_parser& allow_unregistered() { return *this; }
        parsed_options run() const
        {
            parsed_options out;
            out.description = m_desc;
            for (size_t k = 0; k < m_args.size(); ++k) {
                const std::string& a = m_args[k];
                if (a.rfind("--", 0) != 0)
                    continue;
                std::string key = a.substr(2), val;
                auto eq = key.find('=');
                bool haveVal = false;
                if (eq != std::string::npos) {
                    val = key.substr(eq + 1);
                    key = key.substr(0, eq);
                    haveVal = true;
                } else if (k + 1 < m_args.size() && m_args[k + 1].rfind("--", 0) != 0) {
                    val = m_args[++k];
                    haveVal = true;
                }
                basic_option o;
                o.string_key = key;
                if (haveVal)
                    o.value.push_back(val);
                o.unregistered = !(m_desc && m_desc->find(key));
                out.options.push_back(o);
            }
            return out;
        }

    private:
        std::vector<std::string> m_args;
        const options_description* m_desc = nullptr;
    };

    inline parsed_options parse_config_file(
        std::istream& is, const options_description& desc, bool allow_unregistered = false)
    {
        auto trim = [](std::string s) {
            size_t b = s.find_first_not_of(" \t\r\n");
            if (b == std::string::npos)
                return std::string();
            size_t e = s.find_last_not_of(" \t\r\n");
            return s.substr(b, e - b + 1);
        };
        parsed_options out;
        out.description = &desc;
        std::string line, section;
        while (std::getline(is, line)) {
            auto hash = line.find('#');
            if (hash != std::string::npos)
                line = line.substr(0, hash);
            line = trim(line);
            if (line.empty())
                continue;
            if (line.front() == '[' && line.back() == ']') {
                section = trim(line.substr(1, line.size() - 2));
                continue;
            }
            auto eq = line.find('=');
            if (eq == std::string::npos)
                throw std::runtime_error("the options configuration file contains an invalid line '"
                    + line + "'");
            std::string key = trim(line.substr(0, eq));
            std::string val = trim(line.substr(eq + 1));
            if (!section.empty())
                key = section + "." + key;
            basic_option o;
            o.string_key = key;
            o.value.push_back(val);
            o.unregistered = !desc.find(key);
            if (o.unregistered && !allow_unregistered)
                throw std::runtime_error("unrecognised option '" + key + "'");
            out.options.push_back(o);
        }
        return out;
    }

    class variable_value {
    public:
        variable_value() = default;
        variable_value(std::any v, bool defaulted)
            : m_value(std::move(v))
            , m_defaulted(defaulted)
        {
        }
        template <typename T> const T& as() const { return *std::any_cast<T>(&m_value); }
        bool defaulted() const { return m_defaulted; }
        bool empty() const { return !m_value.has_value(); }

    private:
        std::any m_value;
        bool m_defaulted = false;
    };

    class variables_map : public std::map<std::string, variable_value> {
    public:
        const variable_value& operator[](const std::string& name) const
        {
            static const variable_value none;
            auto it = find(name);
            return it == end() ? none : it->second;
        }
        variable_value& slot(const std::string& name)
        {
            return std::map<std::string, variable_value>::operator[](name);
        }
    };

    inline void store(const parsed_options& parsed, variables_map& vm)
    {
        const options_description* desc = parsed.description;
        for (const auto& o : parsed.options) {
            if (o.unregistered || !desc)
                continue;
            const option_description* d = desc->find(o.string_key);
            if (!d || !d->semantic || o.value.empty())
                continue;
            auto it = vm.find(o.string_key);
            if (it != vm.end() && !it->second.defaulted())
                continue; // first explicit value wins
            vm.slot(o.string_key) = variable_value(d->semantic->parse(o.value.front()), false);
        }
        if (desc) {
            for (const auto& d : desc->all()) {
                if (d.semantic && d.semantic->has_default() && vm.find(d.name) == vm.end())
                    vm.slot(d.name) = variable_value(d.semantic->default_any(), true);
            }
        }
    }

    inline void notify(variables_map&) { }

} // namespace program_options
} // namespace boost
