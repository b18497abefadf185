#ifndef ORACLE_SHIM_RAPIDJSON_ISW_H
#define ORACLE_SHIM_RAPIDJSON_ISW_H
#include <istream>
namespace rapidjson {
class IStreamWrapper {
    std::istream& is_;
public:
    explicit IStreamWrapper(std::istream& is) : is_(is) {}
    std::istream& stream() { return is_; }
};
}
#endif
