// TEST INFRASTRUCTURE (oracle/): stand-in for the bison-generated header that
// /root/reference/src/Core/Configuration.cc:899-925 includes.  bison is absent in this
// image; arithmetic expressions in config values are never used by the scorer driver.
#ifndef ORACLE_STUB_ARITHMETIC_EXPRESSION_PARSER_HH
#define ORACLE_STUB_ARITHMETIC_EXPRESSION_PARSER_HH
#include <string>
namespace Core {
class ArithmeticExpressionParserDriver {
public:
    bool parse(const std::string&, double& result) {
        result = 0.0;
        return false;
    }
    std::string getLastError() const {
        return "arithmetic expressions unsupported in the oracle build";
    }
};
}  // namespace Core
#endif
