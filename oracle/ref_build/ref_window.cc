// TEST INFRASTRUCTURE (oracle/): drives the reference's own, UNMODIFIED sliding-window node -- Signal::DelayNode
// (src/Signal/Delay.cc) through Signal::VectorSequenceConcatenation<f32>, filter
// "signal-vector-f32-sequence-concatenation" -- inside a Flow::Network wired as share/lda.flow wires it
// (lda:in -> window -> lda:out, max-size = length, right = right, margin-condition = present-not-empty,
// expand-timestamp = false; the matrix multiplication that follows is pinned by ref_matmul), compiled by
// oracle/ref_build/Makefile with the reference's flags, to pin the window of oracle/gmm_oracle.c orc_lda and the fixture
// tests/golden/lda_window.npz.  The network is built through Flow::Network's own addNode / addLink calls: the XML network
// parser needs libxml2, which this image lacks (stubs/libxml/parser.h lets the headers parse; nothing XML is built).
//
// usage: ref_window in=<file> out=<file> [margin-policy=copy|zero|...] [margin-condition=present-not-empty|...]
//   in : u32 T, baseDim, length, right; f32 frames[T*baseDim]         (one segment)
//   out: u32 nOut, width; f32 windows[nOut*width]
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

#include <Core/Application.hh>
#include <Flow/Attributes.hh>
#include <Flow/Module.hh>
#include <Flow/Network.hh>
#include <Flow/Vector.hh>
#include <Signal/VectorSequenceConcatenation.hh>

class RefWindowApp : public Core::Application {
    std::map<std::string, std::string> args_;

    std::string arg(const std::string& key, const std::string& dflt = "") const {
        std::map<std::string, std::string>::const_iterator i = args_.find(key);
        return i == args_.end() ? dflt : i->second;
    }

public:
    RefWindowApp() {
        setTitle("ref-window");
        setDefaultLoadConfigurationFile(false);
        setDefaultOutputXmlHeader(false);
    }
    std::string getUsage() const {
        return "ref_window in=<file> out=<file> [margin-policy=copy] [margin-condition=present-not-empty]\n";
    }

    int main(const std::vector<std::string>& arguments) {
        for (size_t i = 0; i < arguments.size(); ++i) {
            size_t eq = arguments[i].find('=');
            if (eq != std::string::npos)
                args_[arguments[i].substr(0, eq)] = arguments[i].substr(eq + 1);
        }
        FILE* f = fopen(arg("in").c_str(), "rb");
        if (!f || arg("out").empty()) {
            std::cerr << getUsage();
            return 2;
        }
        u32 hdr[4];
        if (fread(hdr, 4, 4, f) != 4)
            return 3;
        const u32        T = hdr[0], baseDim = hdr[1], length = hdr[2], right = hdr[3];
        std::vector<f32> frames(size_t(T) * baseDim);
        if (!frames.empty() && fread(frames.data(), 4, frames.size(), f) != frames.size())
            return 4;
        fclose(f);

        Flow::Module::instance();  // registers the datatypes (vector-f32)
        Flow::Network net(select("lda"), false);
        net.setTypeName("lda");
        net.addInput("in");
        net.addOutput("out");
        Signal::VectorSequenceConcatenation<f32>* window =
                new Signal::VectorSequenceConcatenation<f32>(net.select("window"));
        window->setParameter("max-size", Core::form("%u", length));
        window->setParameter("right", Core::form("%u", right));
        window->setParameter("margin-condition", arg("margin-condition", "present-not-empty"));
        window->setParameter("margin-policy", arg("margin-policy", "copy"));
        window->setParameter("expand-timestamp", "false");
        if (!net.addNode(window) || !net.addLink("lda", "in", "window", "") || !net.addLink("window", "", "lda", "out"))
            return 5;

        Core::Ref<Flow::Attributes> atts(new Flow::Attributes());
        atts->set("datatype", Flow::Vector<f32>::type()->name());
        net.putAttributes(0, atts);
        for (u32 t = 0; t < T; ++t) {
            Flow::Vector<f32>* v = new Flow::Vector<f32>(frames.begin() + size_t(t) * baseDim,
                                                         frames.begin() + size_t(t + 1) * baseDim);
            v->setStartTime(0.01 * t);
            v->setEndTime(0.01 * t + 0.025);
            net.putData(0, v);
        }
        net.putData(0, Flow::Data::eos());

        std::vector<f32>                   out;
        u32                                nOut = 0, width = 0;
        Flow::DataPtr<Flow::Vector<f32>> d;
        while (net.getData(0, d)) {
            width = d->size();
            out.insert(out.end(), d->begin(), d->end());
            ++nOut;
        }
        f = fopen(arg("out").c_str(), "wb");
        if (!f)
            return 6;
        fwrite(&nOut, 4, 1, f);
        fwrite(&width, 4, 1, f);
        if (!out.empty())
            fwrite(out.data(), 4, out.size(), f);
        fclose(f);
        return 0;
    }
};

APPLICATION(RefWindowApp)
