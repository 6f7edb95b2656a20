/* Minimal stand-in for <SoapySDR/Device.hpp> (oracle/_ref build only):
 * just the names Cariboulite.hpp mentions. No behaviour. */
#pragma once
#include <map>
#include <string>
#include <vector>
#define SOAPY_SDR_TX 0
#define SOAPY_SDR_RX 1
#define SOAPY_SDR_NOT_SUPPORTED (-5)
namespace SoapySDR {
typedef std::map<std::string, std::string> Kwargs;
typedef std::vector<Kwargs> KwargsList;
class Range {
public:
    Range(void) : _min(0), _max(0), _step(0) {}
    Range(const double mn, const double mx, const double st = 0.0) : _min(mn), _max(mx), _step(st) {}
    double minimum(void) const { return _min; }
    double maximum(void) const { return _max; }
    double step(void) const { return _step; }
private:
    double _min, _max, _step;
};
typedef std::vector<Range> RangeList;
class ArgInfo {
public:
    std::string key, value, name, description, units;
    enum Type { BOOL, INT, FLOAT, STRING } type;
    Range range;
    std::vector<std::string> options, optionNames;
};
typedef std::vector<ArgInfo> ArgInfoList;
class Stream;
class Device {
public:
    virtual ~Device(void) {}
};
}
