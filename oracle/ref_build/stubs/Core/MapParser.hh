// TEST INFRASTRUCTURE (oracle/ref_build): stand-in for the reference's Core/MapParser.hh, which needs libxml2 (absent
// in this image).  The only user on the scorer path is Mc::ScaleUpdate::setMap (src/Mc/Component.cc:31-50),
// which reads an optional XML file of model-combination scales (`scales-file`); the harness never sets that parameter.
#ifndef _CORE_MAP_PARSER_HH
#define _CORE_MAP_PARSER_HH

#include <Core/Application.hh>
#include <Core/Component.hh>

namespace Core {

template<class Map>
class XmlMapDocument {
public:
    XmlMapDocument(const Configuration&, Map&, const char*, const char*, const char*, const char*) {}
    bool parseFile(const char* filename) {
        Application::us()->error("scales-file '%s': XML parsing is not available in this harness (no libxml2)", filename);
        return false;
    }
};

}  // namespace Core

#endif  // _CORE_MAP_PARSER_HH
