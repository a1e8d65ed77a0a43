#include "b200_configuration.h"
#include <algorithm>
#include <cctype>
#include <fstream>

namespace gb200h
{
namespace
{
std::string trim(const std::string& s)
{
    size_t b = 0, e = s.size();
    while (b < e && std::isspace(static_cast<unsigned char>(s[b]))) b++;
    while (e > b && std::isspace(static_cast<unsigned char>(s[e - 1]))) e--;
    return s.substr(b, e - b);
}
}  // namespace

int B200Configuration::load_string(const std::string& text)
{
    std::istringstream in(text);
    std::string line;
    int n = 0;
    while (std::getline(in, line))
        {
            std::string l = trim(line);
            if (l.empty() || l[0] == ';' || l[0] == '#' || l[0] == '[') continue;  // comment or section header
            const size_t eq = l.find_first_of("=:");
            if (eq == std::string::npos) continue;
            std::string name = trim(l.substr(0, eq));
            std::string value = l.substr(eq + 1);
            // inline comment: ';' preceded by white space (inih rule used by the reference's INIReader)
            for (size_t i = 1; i < value.size(); i++)
                if (value[i] == ';' && std::isspace(static_cast<unsigned char>(value[i - 1])))
                    {
                        value.resize(i);
                        break;
                    }
            value = trim(value);
            if (name.empty()) continue;
            if (!props_.count(name)) n++;
            props_[name] = value;
        }
    return n;
}

int B200Configuration::load_file(const std::string& path)
{
    std::ifstream f(path);
    if (!f.is_open()) return -1;
    std::stringstream ss;
    ss << f.rdbuf();
    return load_string(ss.str());
}

std::string B200Configuration::property(std::string property_name, std::string default_value) const
{
    auto it = props_.find(property_name);
    return it == props_.end() ? default_value : it->second;
}

bool B200Configuration::property(std::string property_name, bool default_value) const
{
    auto it = props_.find(property_name);
    if (it == props_.end() || it->second.empty()) return default_value;
    std::string v = it->second;
    std::transform(v.begin(), v.end(), v.begin(), [](unsigned char c) { return std::tolower(c); });
    if (v == "true" || v == "1") return true;
    if (v == "false" || v == "0") return false;
    return default_value;
}
}  // namespace gb200h
