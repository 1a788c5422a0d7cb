// TEST INFRASTRUCTURE (oracle/): replaces /root/reference/src/Core/Dependency.cc, whose
// DependencySet::read needs libxml2 (absent here).  Only the symbols other Core TUs link
// against are provided; none is on the scoring path.
#include <Core/Dependency.hh>
#include <Core/StringUtilities.hh>

namespace Core {
const std::string DependencySet::nameSeparator_(".");

void Dependency::setValue(const std::string& v) {
    value_ = v;
}
void DependencySet::add(const std::string& name, const Dependency& d) {
    dependencies_[name] = d;
}
void DependencySet::add(const std::string& name, const std::string& value) {
    add(name, Dependency(value));
}
void DependencySet::add(const std::string& name, const double value) {
    add(name, Dependency(Core::form("%f", value)));
}
bool DependencySet::read(const Configuration&, const std::string&) {
    return false;
}
}  // namespace Core
