// TEST INFRASTRUCTURE (oracle/): shared by the harnesses around the UNMODIFIED reference classes (ref_driver.cc,
// ref_estimator.cc): standard headers, the access-specifier switch, raw-file helpers and the construction of a
// Mm::MixtureSet from a raw "GMS1" model file (format: rasr_b200/gms.py).
#pragma once
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <thread>
#include <vector>
// every std header the reference headers pull in must be seen BEFORE the access-specifier defines
#include <bitset>
#include <complex>
#include <deque>
#include <functional>
#include <iomanip>
#include <list>
#include <locale>
#include <memory>
#include <mutex>
#include <numeric>
#include <set>
#include <sstream>
#include <stack>
#include <unordered_map>
#include <unordered_set>
#include <ext/hash_map>
#include <ext/hash_set>

// The quantized tables and quantizedScore() are private in the reference class; a test harness may
// look inside.  Access specifiers do not change object layout with g++.
#define private public
#define protected public
#include <Core/Application.hh>
#include <Mm/BatchFeatureScorer.hh>
#include <Mm/GaussDiagonalMaximumFeatureScorer.hh>
#include <Mm/MixtureSet.hh>
#include <Mm/SimdFeatureScorer.hh>
#ifdef REF_WITH_ESTIMATORS
#include <Mm/MixtureSetEstimator.hh>
#endif
#undef private
#undef protected

namespace {

template<class T>
bool readArray(FILE* f, std::vector<T>& v, size_t n) {
    v.resize(n);
    return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}

template<class T>
void writeArray(const std::string& path, const std::vector<T>& v) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) {
        fprintf(stderr, "cannot write %s\n", path.c_str());
        exit(2);
    }
    if (!v.empty())
        fwrite(v.data(), sizeof(T), v.size(), f);
    fclose(f);
}

struct RawModel {
    u32              D, M, nDens, nMeans, nCov;
    std::vector<u32> mixOffsets;  // M+1
    std::vector<u32> mixDensity;  // total
    std::vector<f64> mixLogWeight;
    std::vector<u32> densMean, densCov;
    std::vector<f32> means;  // nMeans x D
    std::vector<f32> vars;   // nCov x D
};

bool loadRawModel(const std::string& path, RawModel& r) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f)
        return false;
    char magic[4];
    u32  hdr[5];
    bool ok = fread(magic, 1, 4, f) == 4 && memcmp(magic, "GMS1", 4) == 0 && fread(hdr, 4, 5, f) == 5;
    if (ok) {
        r.D      = hdr[0];
        r.M      = hdr[1];
        r.nDens  = hdr[2];
        r.nMeans = hdr[3];
        r.nCov   = hdr[4];
        ok       = readArray(f, r.mixOffsets, r.M + 1);
    }
    if (ok) {
        size_t total = r.mixOffsets[r.M];
        ok = readArray(f, r.mixDensity, total) && readArray(f, r.mixLogWeight, total) &&
             readArray(f, r.densMean, r.nDens) && readArray(f, r.densCov, r.nDens) &&
             readArray(f, r.means, size_t(r.nMeans) * r.D) && readArray(f, r.vars, size_t(r.nCov) * r.D);
    }
    fclose(f);
    return ok;
}

// Model containers: MixtureSet.hh:66-96,160-204, Mixture.cc:56-61 (addLogDensity), GaussDensity.hh:163-170.
Core::Ref<Mm::MixtureSet> buildMixtureSet(const RawModel& r) {
    Core::Ref<Mm::MixtureSet> ms(new Mm::MixtureSet);
    ms->setDimension(r.D);
    for (u32 i = 0; i < r.nMeans; ++i) {
        Mm::Mean* mean = new Mm::Mean(r.D);
        std::copy(r.means.begin() + size_t(i) * r.D, r.means.begin() + size_t(i + 1) * r.D, mean->begin());
        ms->addMean(i, mean);
    }
    for (u32 c = 0; c < r.nCov; ++c) {
        std::vector<Mm::VarianceType> v(r.vars.begin() + size_t(c) * r.D, r.vars.begin() + size_t(c + 1) * r.D);
        ms->addCovariance(c, new Mm::DiagonalCovariance(v));
    }
    for (u32 d = 0; d < r.nDens; ++d)
        ms->addDensity(d, new Mm::GaussDensity(r.densMean[d], r.densCov[d]));
    for (u32 m = 0; m < r.M; ++m) {
        Mm::Mixture* mix = new Mm::Mixture;
        for (u32 k = r.mixOffsets[m]; k < r.mixOffsets[m + 1]; ++k)
            mix->addLogDensity(r.mixDensity[k], r.mixLogWeight[k]);
        ms->addMixture(m, mix);
    }
    return ms;
}

}  // namespace
