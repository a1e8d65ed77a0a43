// .conf reader for the standalone host layer: the same `name=value` files GNSS-SDR's FileConfiguration reads
// (src/core/receiver/file_configuration.cc: one [GNSS-SDR] section, `;` comments), the same typed property()
// look-ups with defaults (src/core/interfaces/configuration_interface.h:44-59) and the same conversion rule:
// a value that does not parse as the requested type yields the default (src/core/libs/string_converter.cc).
#ifndef GB200_CONFIGURATION_H
#define GB200_CONFIGURATION_H

#include "gnss_sdr_compat.h"
#include <map>
#include <sstream>
#include <string>

namespace gb200h
{
class B200Configuration : public ConfigurationInterface
{
public:
    B200Configuration() = default;
    // parse the text of a .conf file; returns the number of properties read
    int load_string(const std::string& text);
    // read a .conf file; returns the number of properties read, -1 when the file cannot be opened
    int load_file(const std::string& path);
    bool has(const std::string& name) const { return props_.count(name) != 0; }

    std::string property(std::string property_name, std::string default_value) const override;
    bool property(std::string property_name, bool default_value) const override;
    int64_t property(std::string property_name, int64_t default_value) const override { return number(property_name, default_value); }
    uint64_t property(std::string property_name, uint64_t default_value) const override { return number(property_name, default_value); }
    int32_t property(std::string property_name, int32_t default_value) const override { return number(property_name, default_value); }
    uint32_t property(std::string property_name, uint32_t default_value) const override { return number(property_name, default_value); }
    int16_t property(std::string property_name, int16_t default_value) const override { return number(property_name, default_value); }
    uint16_t property(std::string property_name, uint16_t default_value) const override { return number(property_name, default_value); }
    float property(std::string property_name, float default_value) const override { return number(property_name, default_value); }
    double property(std::string property_name, double default_value) const override { return number(property_name, default_value); }
    void set_property(std::string property_name, std::string value) override { props_[property_name] = value; }

private:
    template <typename T>
    T number(const std::string& name, T def) const
    {
        auto it = props_.find(name);
        if (it == props_.end() || it->second.empty()) return def;
        std::stringstream ss(it->second);
        T v;
        ss >> v;
        return ss.fail() ? def : v;
    }
    std::map<std::string, std::string> props_;
};
}  // namespace gb200h
#endif
