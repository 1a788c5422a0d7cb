/* TEST INFRASTRUCTURE -- NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's GMM emission-scoring algorithms (3dimaging/rasr, src/Mm), in
 * plain C, used only as the checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs.  The product (rasr_b200/) never imports, links or executes this file.
 *
 * Parity status: PINNED against the reference itself.  The reference ships no golden vector for any
 * Mm scorer, so the pin is the unmodified reference classes compiled by
 * oracle/ref_build/Makefile into oracle/_ref/ref_scorer and run in the build container:
 * tests/test_oracle_vs_ref.py compares every function below with it, and tests/golden/ holds outputs
 * of that binary (generator: tests/golden/make_golden.py).
 *
 * The alignment consumers (orc_viterbi_align, orc_baum_welch_align) are pinned the same way: against the
 * unmodified Search::Aligner compiled into oracle/_ref/ref_aligner; see the notes above those functions.
 *
 * Compile WITHOUT -ffast-math: every operation is written in the order the reference binary executes
 * it (its -ffast-math build was inspected by disassembly where the source leaves the order open).
 *
 * Each function cites the reference file:line it follows (paths relative to /root/reference/src/Mm).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <xmmintrin.h>

typedef struct {
    uint32_t        D, M, nDens, nMeans, nCov;
    const uint32_t* mixOffsets;   /* [M+1] */
    const uint32_t* mixDensity;   /* [G]   density index per (mixture, k) */
    const double*   mixLogWeight; /* [G]   */
    const uint32_t* densMean;     /* [nDens] */
    const uint32_t* densCov;      /* [nDens] */
    const float*    means;        /* [nMeans x D] */
    const float*    variances;    /* [nCov x D]   */
} orc_model;

/* ---------------------------------------------------------------------------------------------
 * Table preparation
 * ------------------------------------------------------------------------------------------- */

/* Utilities.hh:86-91 inverseSquareRoot<T>: (T)1 / (T)sqrt(x).  Under the reference's flags
 * (-O2 -ffast-math -msse3, config/proc-x86_64.make:13-29) g++ emits rsqrtss plus one Newton step:
 *     r = rsqrtss(x); out = ((x*r)*r + -3.0f) * (r * -0.5f)
 * (objdump of CovarianceFeatureScorerElement.o, calculateInverseSquareRootDiagonal).  exact != 0
 * gives the IEEE value instead (what a non-fast-math build of the reference would store). */
void orc_inverse_sqrt(const float* var, float* isr, size_t n, int exact) {
    for (size_t i = 0; i < n; ++i) {
        float x = var[i];
        if (exact) {
            isr[i] = 1.0f / sqrtf(x);
        }
        else {
            float r = _mm_cvtss_f32(_mm_rsqrt_ss(_mm_set_ss(x)));
            float t = x * r;
            t       = t * r;
            t       = t + -3.0f;
            isr[i]  = t * (r * -0.5f);
        }
    }
}

/* Utilities.hh:52-58,72-76 gaussLogNormFactor: D*log(2*pi) + sum_d log|var_d| in f64, stored f32
 * (CovarianceFeatureScorerElement.cc:36-39). */
float orc_log_norm_factor(const float* var, uint32_t D) {
    double s = 0.0;
    for (uint32_t d = 0; d < D; ++d)
        s += log(fabs((double)var[d]));
    return (float)((double)D * log((double)2 * M_PI) + s);
}

/* CovarianceFeatureScorerElement.cc:20-51 (operator= then scale(factor)) for every covariance.
 * GaussDiagonalMaximumFeatureScorer.cc:52,82-85: factor = sqrt(gaussian-scale) (1 by default). */
void orc_covariance_tables(const orc_model* m, float factor, int exactRsqrt, float* isr /*[nCov x D]*/,
                           float* logNorm /*[nCov]*/) {
    orc_inverse_sqrt(m->variances, isr, (size_t)m->nCov * m->D, exactRsqrt);
    for (uint32_t c = 0; c < m->nCov; ++c) {
        logNorm[c] = orc_log_norm_factor(m->variances + (size_t)c * m->D, m->D);
        /* scale(): isr *= factor ; logNorm *= factor*factor   (:45-51) */
        for (uint32_t d = 0; d < m->D; ++d)
            isr[(size_t)c * m->D + d] = isr[(size_t)c * m->D + d] * factor;
        logNorm[c] *= factor * factor;
    }
}

/* MixtureFeatureScorerElement.cc:21-33: w2 = f32(-2 * logw) then *= mixture-weight-scale. */
void orc_minus2_log_weights(const orc_model* m, float mixtureWeightScale, float* w2 /*[G]*/) {
    size_t G = m->mixOffsets[m->M];
    for (size_t i = 0; i < G; ++i) {
        float v = (float)(-2 * m->mixLogWeight[i]);
        w2[i]   = v * mixtureWeightScale;
    }
}

/* ---------------------------------------------------------------------------------------------
 * Float scorers
 * ------------------------------------------------------------------------------------------- */

/* GaussDiagonalMaximumFeatureScorer.cc:144-218, the __SSE3__ branch: four lane sums over
 * d < 4*floor(D/4), hadd -> (l0+l1, l2+l3), result = (l0+l1) + (l2+l3), then the scalar tail. */
static float orc_distance(const float* x, const float* mean, const float* isr, uint32_t D) {
    float    lane[4] = {0.f, 0.f, 0.f, 0.f};
    uint32_t eff     = D & ~3u, d = 0;
    for (; d < eff; d += 4) {
        for (int l = 0; l < 4; ++l) {
            float df = (mean[d + l] - x[d + l]) * isr[d + l];
            lane[l] += df * df;
        }
    }
    float b0     = lane[0] + lane[1];
    float b1     = lane[2] + lane[3];
    float result = 0.f;
    result += b0 + b1;
    for (; d < D; ++d) {
        float df = (mean[d] - x[d]) * isr[d];
        result += df * df;
    }
    return result;
}

/* GaussDiagonalMaximumFeatureScorer.cc:116-142 calculateScoreAndDensity for all mixtures of all
 * frames.  bestScore is an f32 that is compared (promoted) against the f64 candidate and assigned
 * (rounded) from it; the first strict improvement wins. */
void orc_score_max(const orc_model* m, const float* isr, const float* logNorm, const float* w2, const float* x,
                   size_t T, float* score /*[T x M]*/, uint32_t* best /*[T x M]*/) {
#pragma omp parallel for schedule(static)
    for (long t = 0; t < (long)T; ++t) {
        const float* xt = x + (size_t)t * m->D;
        for (uint32_t mix = 0; mix < m->M; ++mix) {
            float  bestScore   = FLT_MAX;    /* Core::Type<Score>::max */
            size_t bestDensity = (size_t)-1; /* Core::Type<size_t>::max */
            for (uint32_t k = m->mixOffsets[mix]; k < m->mixOffsets[mix + 1]; ++k) {
                uint32_t dns = m->mixDensity[k];
                uint32_t cov = m->densCov[dns];
                double   s   = (double)w2[k] + (double)logNorm[cov] +
                           (double)orc_distance(xt, m->means + (size_t)m->densMean[dns] * m->D,
                                                isr + (size_t)cov * m->D, m->D);
                if ((double)bestScore > s) {
                    bestScore   = (float)s;
                    bestDensity = k - m->mixOffsets[mix];
                }
            }
            score[(size_t)t * m->M + mix] = (float)(0.5 * bestScore);
            if (best)
                best[(size_t)t * m->M + mix] = (uint32_t)bestDensity;
        }
    }
}

/* GaussDiagonalMaximumFeatureScorer.cc:239-289 (GaussDiagonalSumFeatureScorer): all f32;
 * s_k = 0.5*(w2 + L + dist); best = first min; score = best - log(sum_k exp(best - s_k)).
 * `0.5 * score` is a double product rounded to f32 (:258); std::exp/std::log on f32 are expf/logf.
 * posterior != NULL additionally returns exp(score - s_k) per density slot (:291-298), [T x G]. */
void orc_score_sum(const orc_model* m, const float* isr, const float* logNorm, const float* w2, const float* x,
                   size_t T, float* score, uint32_t* best, float* posterior) {
    size_t G = m->mixOffsets[m->M];
#pragma omp parallel for schedule(static)
    for (long t = 0; t < (long)T; ++t) {
        const float* xt = x + (size_t)t * m->D;
        float*       sc = (float*)malloc(sizeof(float) * (G ? G : 1));
        for (uint32_t mix = 0; mix < m->M; ++mix) {
            uint32_t a = m->mixOffsets[mix], b = m->mixOffsets[mix + 1];
            for (uint32_t k = a; k < b; ++k) {
                uint32_t dns = m->mixDensity[k];
                uint32_t cov = m->densCov[dns];
                float    s   = w2[k] + logNorm[cov] +
                          orc_distance(xt, m->means + (size_t)m->densMean[dns] * m->D, isr + (size_t)cov * m->D, m->D);
                sc[k] = (float)(0.5 * s);
            }
            float  bestScore   = FLT_MAX;
            size_t bestDensity = (size_t)-1;
            for (uint32_t k = a; k < b; ++k) {
                if (bestScore > sc[k]) {
                    bestScore   = sc[k];
                    bestDensity = k - a;
                }
            }
            float sumExp = 0.f;
            for (uint32_t k = a; k < b; ++k)
                sumExp += expf(bestScore - sc[k]);
            float r                       = bestScore - logf(sumExp);
            score[(size_t)t * m->M + mix] = r;
            if (best)
                best[(size_t)t * m->M + mix] = (uint32_t)bestDensity;
            if (posterior)
                for (uint32_t k = a; k < b; ++k)
                    posterior[(size_t)t * G + k] = expf(r - sc[k]);
        }
        free(sc);
    }
}

/* ---------------------------------------------------------------------------------------------
 * Quantized SIMD scorer
 * ------------------------------------------------------------------------------------------- */

/* Utilities.hh:179-191 quantize<f32,u8>: clip((int)round(x) + 128, 0, 255); round() is half away
 * from zero.  (int) of an out-of-range float is the x86 cvttss2si "integer indefinite" 0x80000000,
 * which the clip turns into 0 after the +128 (it stays negative). */
static uint8_t orc_quantize(float x) {
    int r = _mm_cvttss_si32(_mm_set_ss(roundf(x)));
    int v = (int)((unsigned)r + 128u); /* wraps like the reference's int add */
    if (v < 0)
        v = 0;
    if (v > 255)
        v = 255;
    return (uint8_t)v;
}

/* IntelOptimization.hh:55-57 optimalVectorSize with BlockSize = 16 (ENABLE_SSE2). */
uint32_t orc_simd_padded_dimension(uint32_t D) {
    return ((D + 15u) / 16u) * 16u;
}

/* IntelOptimization.cc:50-67 multiplyAndQuantize: r[d] = Q(x[d] * y[d]), zero padded to Dpad. */
static void orc_multiply_and_quantize(const float* x, const float* y, uint32_t D, uint32_t Dpad, uint8_t* r) {
    uint32_t d = 0;
    for (; d < D; ++d)
        r[d] = orc_quantize(x[d] * y[d]);
    for (; d < Dpad; ++d)
        r[d] = 0;
}

/* SimdFeatureScorer.cc:64-133 init/getScaling/quantizationScalingFactor/buildMixtureTable and
 * IntelOptimization.cc:39-48 createDensityElement.
 * Outputs: isrS [nCov x D] (isr * scaling), logNormS [nCov] (L * scaling^2), scalingSquared,
 * qmean [G x Dpad], qconst [G], qcov [G] -- one entry per (mixture, density slot), untied. */
void orc_simd_prepare(const orc_model* m, int exactRsqrt, float* isrS, float* logNormS, float* scalingSquared,
                      uint8_t* qmean, int32_t* qconst, uint32_t* qcov) {
    const uint32_t D = m->D, Dpad = orc_simd_padded_dimension(D);
    /* init(): covarianceTable_[i] = covariance (unscaled)                       (:64-67) */
    orc_covariance_tables(m, 1.0f, exactRsqrt, isrS, logNormS);
    /* getScaling(): min/max over all densities of mean*isr                      (:106-126) */
    float minMean = FLT_MAX, maxMean = -FLT_MAX; /* Core::Type<f32>::max / ::min */
    for (uint32_t i = 0; i < m->nDens; ++i) {
        const float* mean = m->means + (size_t)m->densMean[i] * D;
        const float* isr  = isrS + (size_t)m->densCov[i] * D;
        for (uint32_t d = 0; d < D; ++d) {
            float dividedMean = mean[d] * isr[d];
            minMean           = dividedMean < minMean ? dividedMean : minMean;
            maxMean           = maxMean < dividedMean ? dividedMean : maxMean;
        }
    }
    /* quantizationScalingFactor(): (f32)255 / (1.25 * intervalSize), double division  (:128-133) */
    float a            = fabsf(minMean), b = fabsf(maxMean);
    float intervalSize = 2 * (a < b ? b : a);
    float scaling      = (float)((double)(float)255 / (1.25 * (double)intervalSize));
    *scalingSquared    = scaling * scaling; /* (:72) */
    /* covarianceTable_[i].scale(scaling)                                        (:75-76) */
    for (uint32_t c = 0; c < m->nCov; ++c) {
        for (uint32_t d = 0; d < D; ++d)
            isrS[(size_t)c * D + d] = isrS[(size_t)c * D + d] * scaling;
        logNormS[c] *= scaling * scaling;
    }
    /* buildMixtureTable()                                                       (:81-104) */
    for (uint32_t mix = 0; mix < m->M; ++mix) {
        for (uint32_t k = m->mixOffsets[mix]; k < m->mixOffsets[mix + 1]; ++k) {
            uint32_t dns = m->mixDensity[k];
            uint32_t cov = m->densCov[dns];
            qcov[k]      = cov;
            /* Weight(f64) scaledMinus2LogWeight = scalingSquared_ * -2 * logWeight: (f32 * int) * f64 */
            double scaledMinus2LogWeight = (double)(*scalingSquared * -2) * m->mixLogWeight[k];
            /* passed as Score (f32) into createDensityElement, added to f32 L, truncated to s32 */
            float w = (float)scaledMinus2LogWeight;
            orc_multiply_and_quantize(m->means + (size_t)m->densMean[dns] * D, isrS + (size_t)cov * D, D, Dpad,
                                      qmean + (size_t)k * Dpad);
            qconst[k] = _mm_cvttss_si32(_mm_set_ss(w + logNormS[cov]));
        }
    }
}

/* BatchFeatureScorer.cc:319-376 BatchIntFeatureScorer::quantizationScale / init (`batch-diagonal-maximum-int`):
 * the same u8 grid as the SIMD scorer -- variance_ = isr * scale, means and frames through quantize(a * b) -- but
 * its own constants: c = (s32)(logNormFactor - scale_ * logWeight) with logNormFactor = L * scale^2 (f32),
 * scale_ = 2 * scale^2 (f32), logWeight f64, evaluated in double and truncated (cvttsd2si: 0x80000000 when out of
 * range -- a zero-weight density then WINS every minimum, as it does in the reference).  Pooled covariance only
 * (:340-342 criticalError): returns -1 otherwise.
 * Outputs: isrS [D], scale2 (= scale_, the divisor of the final score, :468), qmean [G x Dpad], qconst [G]. */
int orc_batchint_prepare(const orc_model* m, int exactRsqrt, float* isrS, float* scale2, uint8_t* qmean,
                         int32_t* qconst) {
    if (m->nCov != 1)
        return -1;
    const uint32_t D = m->D, Dpad = orc_simd_padded_dimension(D); /* BlockSize = 16 (:293) */
    float          logNorm;
    orc_covariance_tables(m, 1.0f, exactRsqrt, isrS, &logNorm);
    float minMean = FLT_MAX, maxMean = -FLT_MAX;
    for (uint32_t i = 0; i < m->nDens; ++i) {
        const float* mean = m->means + (size_t)m->densMean[i] * D;
        for (uint32_t d = 0; d < D; ++d) {
            float dividedMean = mean[d] * isrS[d];
            minMean           = dividedMean < minMean ? dividedMean : minMean;
            maxMean           = maxMean < dividedMean ? dividedMean : maxMean;
        }
    }
    float a = fabsf(minMean), b = fabsf(maxMean);
    float intervalSize = 2 * (a < b ? b : a);
    float scale        = (float)((double)(float)255 / (1.25 * (double)intervalSize));
    float scaleSquared = scale * scale;
    *scale2            = (float)(2.0 * (double)scaleSquared);
    for (uint32_t d = 0; d < D; ++d)
        isrS[d] = isrS[d] * scale;
    const float logNormFactor = logNorm * scaleSquared;
    for (uint32_t mix = 0; mix < m->M; ++mix) {
        for (uint32_t k = m->mixOffsets[mix]; k < m->mixOffsets[mix + 1]; ++k) {
            uint32_t dns = m->mixDensity[k];
            orc_multiply_and_quantize(m->means + (size_t)m->densMean[dns] * D, isrS, D, Dpad, qmean + (size_t)k * Dpad);
            qconst[k] = _mm_cvttsd_si32(_mm_set_sd((double)logNormFactor - (double)*scale2 * m->mixLogWeight[k]));
        }
    }
    return 0;
}

/* BatchFeatureScorer.cc:468: *result = static_cast<f32>(best) / scale_  (f32 division) */
void orc_batchint_final_score(const int32_t* q, size_t n, float scale2, float* score) {
    for (size_t i = 0; i < n; ++i) {
        volatile float num = (float)q[i]; /* keeps the division a division under any optimisation flags */
        score[i]           = num / scale2;
    }
}

/* SimdFeatureScorer.cc:22-35 Context ctor: the frame is quantized once per covariance.
 * qx [T x nCov x Dpad]. */
void orc_simd_quantize_features(const float* x, size_t T, uint32_t D, uint32_t nCov, const float* isrS,
                                uint8_t* qx) {
    const uint32_t Dpad = orc_simd_padded_dimension(D);
#pragma omp parallel for schedule(static)
    for (long t = 0; t < (long)T; ++t)
        for (uint32_t c = 0; c < nCov; ++c)
            orc_multiply_and_quantize(x + (size_t)t * D, isrS + (size_t)c * D, D, Dpad,
                                      qx + ((size_t)t * nCov + c) * Dpad);
}

/* SSE2CodeGenerator.cc:324-374 / scalar restatement IntelOptimization.cc:71-113:
 * sum over all Dpad bytes of (mean - feature)^2 in s32. */
static int32_t orc_u8_distance(const uint8_t* a, const uint8_t* b, uint32_t Dpad) {
    int32_t s = 0;
    for (uint32_t d = 0; d < Dpad; ++d) {
        int32_t df = (int32_t)a[d] - (int32_t)b[d];
        s += df * df;
    }
    return s;
}

/* SimdFeatureScorer.cc:158-176 quantizedScore: q = min_k (const_k + dist_k), first strict minimum. */
void orc_simd_score(uint32_t M, const uint32_t* mixOffsets, uint32_t Dpad, uint32_t nCov, const uint8_t* qmean,
                    const int32_t* qconst, const uint32_t* qcov, const uint8_t* qx, size_t T,
                    int32_t* q /*[T x M]*/, uint32_t* best /*[T x M]*/) {
#pragma omp parallel for schedule(static)
    for (long t = 0; t < (long)T; ++t) {
        for (uint32_t mix = 0; mix < M; ++mix) {
            int32_t minScore    = INT32_MAX;
            size_t  bestDensity = (size_t)-1;
            for (uint32_t k = mixOffsets[mix]; k < mixOffsets[mix + 1]; ++k) {
                /* int + int; wraps like the reference build when const is the 0x80000000 indefinite */
                int32_t s = (int32_t)((uint32_t)qconst[k] +
                                      (uint32_t)orc_u8_distance(qmean + (size_t)k * Dpad,
                                                                qx + ((size_t)t * nCov + qcov[k]) * Dpad, Dpad));
                if (s < minScore) {
                    minScore    = s;
                    bestDensity = k - mixOffsets[mix];
                }
            }
            q[(size_t)t * M + mix] = minScore;
            if (best)
                best[(size_t)t * M + mix] = (uint32_t)bestDensity;
        }
    }
}

/* SimdFeatureScorer.cc:135-145: score = f32( 0.5 * q / scalingSquared_ ), double arithmetic. */
void orc_simd_final_score(const int32_t* q, size_t n, float scalingSquared, float* score) {
    for (size_t i = 0; i < n; ++i)
        score[i] = (float)(0.5 * q[i] / (double)scalingSquared);
}

/* ---------------------------------------------------------------------------------------------
 * Viterbi alignment over a score block (the consumer of the scores)
 *
 * PINNED against the reference itself: oracle/ref_build/Makefile compiles the unmodified Search/Aligner.cc with the
 * reference's Fsa library, Speech/Alignment.cc and Am/AcousticModel.cc into oracle/_ref/ref_aligner (driver:
 * oracle/ref_build/ref_aligner.cc -- a table-backed acoustic model and per-frame scorers stand in for the model files;
 * libxml2 being absent, stubs/libxml/parser.h names the types the Core headers mention and no XML code is built).
 * tests/test_oracle_vs_ref.py compares this function with it on fresh random automata -- same emission at every
 * frame, same f32 score, exact ties included -- and tests/golden/aln_*.npz hold that binary's answers
 * (generator: tests/golden/make_align_golden.py).  The reference holds no golden alignment of its own.
 * Paths relative to /root/reference/src/Search.
 * ------------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t state;
    float    score;
    uint32_t trace;    /* index of the predecessor hypothesis */
    uint32_t emission; /* Aligner.cc:205 stores the arc label; the emission index stands in for it here */
} orc_state_hyp;

/* One segment.  Automaton as in include/gmm_b200.h: arcs of state s are arcOffsets[s]..arcOffsets[s+1] in Fsa
 * order, finalWeight[s] = +inf marks a non-final state.  Returns the best final score (FLT_MAX and a path of
 * 0xFFFFFFFF when no final state is reached, Aligner.cc:417-420). */
float orc_viterbi_align(uint32_t T, const float* scores, size_t ld, uint32_t S, const uint32_t* arcOffsets,
                        const uint32_t* arcTarget, const uint32_t* arcEmission, const float* arcWeight,
                        uint32_t initial, const float* finalWeight, uint32_t* pathEmission, uint32_t* pathState) {
    size_t         cap = 1024, n = 0;
    orc_state_hyp* hyp = (orc_state_hyp*)malloc(cap * sizeof(orc_state_hyp));
    size_t*        hypForState = (size_t*)calloc(S ? S : 1, sizeof(size_t));
    /* Aligner.cc:176-181 addStartupHypothesis: score 0 at the initial state */
    hyp[n++]             = (orc_state_hyp){initial, 0.0f, 0xFFFFFFFFu, 0xFFFFFFFFu};
    hypForState[initial] = 0;
    size_t firstCur      = 0;
    for (uint32_t t = 0; t < T; ++t) {
        /* Aligner.cc:190-239 expand */
        const float* row       = scores + (size_t)t * ld;
        size_t       firstPrev = firstCur;
        firstCur               = n;
        for (size_t shi = firstPrev; shi < firstCur; ++shi) {
            const uint32_t st    = hyp[shi].state;
            const float    score = hyp[shi].score;
            for (uint32_t a = arcOffsets[st]; a < arcOffsets[st + 1]; ++a) {
                const float emissionScore   = row[arcEmission[a]];
                const float transitionScore = arcWeight[a];
                volatile float partial      = score + transitionScore; /* Aligner.cc:207, left to right in f32 */
                const float    sco          = partial + emissionScore;
                /* Aligner.cc:132-158 activateOrUpdateStateHypothesisViterbi */
                const uint32_t target = arcTarget[a];
                size_t         k      = hypForState[target];
                if (k >= n || k < firstCur || hyp[k].state != target) {
                    hypForState[target] = n;
                    if (n == cap) {
                        cap *= 2;
                        hyp = (orc_state_hyp*)realloc(hyp, cap * sizeof(orc_state_hyp));
                    }
                    hyp[n++] = (orc_state_hyp){target, sco, (uint32_t)shi, arcEmission[a]};
                }
                else if (hyp[k].score >= sco) {
                    hyp[k].score    = sco;
                    hyp[k].trace    = (uint32_t)shi;
                    hyp[k].emission = arcEmission[a];
                }
            }
        }
    }
    /* Aligner.cc:398-411 getAlignmentFsaViterbi: best final hypothesis of the last frame, strict < */
    size_t bestHyp   = (size_t)-1;
    float  bestScore = FLT_MAX;
    for (size_t shi = firstCur; shi < n; ++shi) {
        const float fw = finalWeight[hyp[shi].state];
        if (isinf(fw) && fw > 0)
            continue; /* not final */
        const float sc = fw + hyp[shi].score;
        if (sc < bestScore) {
            bestScore = sc;
            bestHyp   = shi;
        }
    }
    if (bestScore == FLT_MAX) {
        for (uint32_t t = 0; t < T; ++t) {
            pathEmission[t] = 0xFFFFFFFFu;
            if (pathState)
                pathState[t] = 0xFFFFFFFFu;
        }
    }
    else {
        /* Aligner.cc:425-434 trace back */
        size_t shi = bestHyp;
        for (uint32_t t = T; t > 0; --t) {
            pathEmission[t - 1] = hyp[shi].emission;
            if (pathState)
                pathState[t - 1] = hyp[shi].state;
            shi = hyp[shi].trace;
        }
    }
    free(hyp);
    free(hypForState);
    return bestScore;
}

/* ---------------------------------------------------------------------------------------------
 * Training E-step: best density of the aligned mixture + accumulation.
 *
 * PINNED: oracle/ref_build/ref_estimator.cc links the unmodified estimator classes; tests/test_oracle_vs_ref.py and
 * tests/golden/est_*.npz compare the accumulators below with the MIXSET files it writes, bit for bit in viterbi mode.
 * The best density and the posteriors are the pinned scorer functions above restricted to one mixture.  Follows, in
 * frame order and in f64 like the reference:
 *   Speech/KeyedEstimator.cc:108-121 processAlignedFeature
 *   AbstractMixtureSetEstimator.cc:117-147 accumulate (viterbi / posterior-weighted with weightThreshold)
 *   MixtureEstimator.cc:86-97, GaussDensityEstimator.cc:22-32, VectorAccumulator.hh:60-67,
 *   Utilities.hh:110-146 (sum + y, sum + y*y, sum + w*y; the source's sum + w*y*y is evaluated as sum + w*(y*y) by the
 *   reference build, -ffast-math: checked bit for bit against oracle/_ref/ref_estimator's accumulator files)
 * sumScorer = 0: GaussDiagonalMaximumFeatureScorer's best density; 1: GaussDiagonalSumFeatureScorer's.
 * viterbi = 0 needs sumScorer = 1 (only the sum scorer has density posteriors).
 * mixture[t] = 0xFFFFFFFF skips frame t; weight may be NULL (the unweighted overloads).
 * Accumulators: meanSum [nMeans x D], meanW [nMeans], covSum [nCov x D], covW [nCov], mixW [G]; best [T] or NULL.
 * ------------------------------------------------------------------------------------------- */
static void orc_accumulate_density(const orc_model* m, uint32_t slot, const float* xt, double w, int weighted,
                                   double* meanSum, double* meanW, double* covSum, double* covW, double* mixW) {
    uint32_t dns = m->mixDensity[slot], mean = m->densMean[dns], cov = m->densCov[dns];
    for (uint32_t d = 0; d < m->D; ++d) {
        double y = (double)xt[d];
        if (weighted) {
            meanSum[(size_t)mean * m->D + d] = meanSum[(size_t)mean * m->D + d] + w * y;
            covSum[(size_t)cov * m->D + d]   = covSum[(size_t)cov * m->D + d] + w * (y * y);
        }
        else {
            meanSum[(size_t)mean * m->D + d] = meanSum[(size_t)mean * m->D + d] + y;
            covSum[(size_t)cov * m->D + d]   = covSum[(size_t)cov * m->D + d] + y * y;
        }
    }
    /* unweighted: ++weight_ (VectorAccumulator.hh:61), ++weights_[i] (MixtureEstimator.cc:88) */
    meanW[mean] += weighted ? w : 1.0;
    covW[cov] += weighted ? w : 1.0;
    mixW[slot] += weighted ? w : 1.0;
}

void orc_accumulate(const orc_model* m, const float* isr, const float* logNorm, const float* w2, const float* x,
                    size_t T, const uint32_t* mixture, const double* weight, int sumScorer, int viterbi,
                    double weightThreshold, double* meanSum, double* meanW, double* covSum, double* covW,
                    double* mixW, uint32_t* best) {
    float* sc = (float*)malloc(sizeof(float) * (m->mixOffsets[m->M] ? m->mixOffsets[m->M] : 1));
    for (size_t t = 0; t < T; ++t) {
        uint32_t mix = mixture[t];
        if (best)
            best[t] = 0xFFFFFFFFu;
        if (mix == 0xFFFFFFFFu)
            continue;
        const float* xt = x + t * m->D;
        uint32_t     a = m->mixOffsets[mix], b = m->mixOffsets[mix + 1];
        float        bestScore   = FLT_MAX;
        size_t       bestDensity = (size_t)-1;
        for (uint32_t k = a; k < b; ++k) {
            uint32_t dns = m->mixDensity[k], cov = m->densCov[dns];
            float    dist =
                    orc_distance(xt, m->means + (size_t)m->densMean[dns] * m->D, isr + (size_t)cov * m->D, m->D);
            if (sumScorer) {
                float s = w2[k] + logNorm[cov] + dist;
                sc[k]   = (float)(0.5 * s);
                if (bestScore > sc[k]) {
                    bestScore   = sc[k];
                    bestDensity = k - a;
                }
            }
            else {
                double s = (double)w2[k] + (double)logNorm[cov] + (double)dist;
                if ((double)bestScore > s) {
                    bestScore   = (float)s;
                    bestDensity = k - a;
                }
            }
        }
        if (best)
            best[t] = (uint32_t)bestDensity;
        double w = weight ? weight[t] : 1.0;
        if (viterbi) {
            if (bestDensity != (size_t)-1)
                orc_accumulate_density(m, a + (uint32_t)bestDensity, xt, w, weight != NULL, meanSum, meanW, covSum,
                                       covW, mixW);
        }
        else if (sumScorer) {
            float sumExp = 0.f;
            for (uint32_t k = a; k < b; ++k)
                sumExp += expf(bestScore - sc[k]);
            float score = bestScore - logf(sumExp);
            for (uint32_t k = a; k < b; ++k) {
                double dnsWeight   = (double)expf(score - sc[k]); /* std::vector<Mm::Weight> of f32 exp results */
                double finalWeight = w * dnsWeight;
                if (finalWeight > weightThreshold)
                    orc_accumulate_density(m, k, xt, finalWeight, 1, meanSum, meanW, covSum, covW, mixW);
            }
        }
    }
    free(sc);
}

/* ---------------------------------------------------------------------------------------------
 * Front end: sliding window + matrix multiplication (lda.flow)
 *
 * The matrix product is PINNED: tests/test_oracle_vs_ref.py compares it with the reference's own
 * Math::Matrix<f32> * Math::Vector<f32> (oracle/_ref/ref_matmul, built by oracle/ref_build/Makefile);
 * tests/golden/lda_matmul.npz holds outputs of that binary.  The window is PINNED too: oracle/_ref/ref_window runs the
 * reference's own Signal::DelayNode (through signal-vector-f32-sequence-concatenation in a Flow::Network wired as
 * share/lda.flow); tests/test_lda_oracle.py compares on fresh random shapes, tests/golden/lda_window.npz holds that
 * binary's windows and, multiplied by ref_matmul, the whole front end's outputs.
 *   Signal/VectorSequenceConcatenation.hh:85-100  window of frames t-left .. t+right, oldest first
 *   Signal/Delay.hh:32-46, Delay.cc:30-31,171-174  margin-condition present-not-empty (lda.flow), margin-policy
 *                                                  copy: a missing frame is the closest existing one
 *   Math/Matrix.hh:482-489, Math/Vector.hh:95-101  result[n] = sum_i a[n][i] * v[i], f32, left to right (the
 *                                                  reference build keeps the scalar loop: mulss / addss)
 * base [T x baseDim]; segments [frameOffsets[s], frameOffsets[s+1]); matrix [D x baseDim*length]; out [T x D].
 * ------------------------------------------------------------------------------------------- */
void orc_matrix_vector(const float* matrix, uint32_t rows, uint32_t cols, const float* v, float* out) {
    for (uint32_t n = 0; n < rows; ++n) {
        float result = 0.f;
        for (uint32_t i = 0; i < cols; ++i)
            result += matrix[(size_t)n * cols + i] * v[i];
        out[n] = result;
    }
}

void orc_lda(const float* base, uint32_t nSeg, const uint32_t* frameOffsets, uint32_t baseDim, uint32_t length,
             uint32_t right, const float* matrix, uint32_t D, float* out) {
    const uint32_t Din = baseDim * length;
    float*         win = (float*)malloc(sizeof(float) * (Din ? Din : 1));
    const long     left = (long)length - 1 - (long)right;
    for (uint32_t s = 0; s < nSeg; ++s) {
        const long b = frameOffsets[s], e = frameOffsets[s + 1];
        for (long t = b; t < e; ++t) {
            for (long r = -left; r <= (long)right; ++r) {
                long f = t + r;
                f      = f < b ? b : f >= e ? e - 1 : f;
                memcpy(win + (size_t)(r + left) * baseDim, base + (size_t)f * baseDim, sizeof(float) * baseDim);
            }
            orc_matrix_vector(matrix, D, Din, win, out + (size_t)t * D);
        }
    }
    free(win);
}

/* ---------------------------------------------------------------------------------------------
 * Baum-Welch alignment (Search::Aligner in mode baum-welch, unpruned, min-weight 0).  PINNED as orc_viterbi_align:
 * against oracle/_ref/ref_aligner = the unmodified aligner with Fsa::posterior64 and Speech::Alignment's combineItems /
 * clipWeights / shiftMinToZeroWeights / expm / normalizeWeights (tests/test_oracle_vs_ref.py, tests/golden/aln_*.npz):
 * state posteriors agree to 1e-5 absolute (the reference rounds arc posteriors to f32 -log values and combines them in
 * the f32 log semiring; this restatement stays in f64), totals to 1e-6 relative.
 *   Search/Aligner.cc:160-187  activateOrUpdateStateHypothesisBaumWelch: the forward potential of a state is
 *                              collected in the log semiring over its incoming arcs; the traceback lattice keeps
 *                              one arc per (frame, automaton arc) with weight transition + emission score
 *   Search/Aligner.cc:438-452  getAlignmentFsaBaumWelch: final weights close the lattice
 *   Search/Aligner.cc:757-772  getAlignmentPosteriorFsa: Fsa::posterior64 -- arc posteriors (f64 forward / backward),
 *                              as -log probabilities
 *   Search/Aligner.cc:693-716  getAlignment: items (time, emission, -log posterior) combined per (time, emission) in
 *                              the log semiring, clipped to >= 0, shifted, exp(-.), normalised per time frame, > 0 kept
 *   Speech/Alignment.cc:455-566 combineItems / clipWeights / shiftMinToZeroWeights / expm / normalizeWeights
 * Here in f64 throughout: weights[t][j] = posterior of emission emList[j] at frame t (0 where none), emList = the
 * distinct emissions of the automaton in order of first use by an arc.  Returns the total score -log p(x | automaton)
 * (FLT_MAX and all-zero weights when no final state is reached: empty alignment, Aligner.cc:734-735).
 * ------------------------------------------------------------------------------------------- */
static double orc_log_add(double a, double b) { /* Fsa::LogSemiring::collect: -log(exp(-a) + exp(-b)) */
    if (isinf(a) && a > 0)
        return b;
    if (isinf(b) && b > 0)
        return a;
    const double m = a < b ? a : b, d = a < b ? b - a : a - b;
    return m - log1p(exp(-d));
}

double orc_baum_welch_align(uint32_t T, const float* scores, size_t ld, uint32_t S, const uint32_t* arcOffsets,
                            const uint32_t* arcTarget, const uint32_t* arcEmission, const float* arcWeight,
                            uint32_t initial, const float* finalWeight, uint32_t nList, const uint32_t* emList,
                            float* weights /* [T x ldW] */, size_t ldW) {
    double* alpha = (double*)malloc(sizeof(double) * (size_t)(T + 1) * (S ? S : 1));
    double* beta  = (double*)malloc(sizeof(double) * 2 * (S ? S : 1));
    double* gamma = (double*)malloc(sizeof(double) * (nList ? nList : 1));
    for (size_t i = 0; i < (size_t)(T + 1) * S; ++i)
        alpha[i] = INFINITY;
    for (uint32_t t = 0; t < T; ++t)
        for (uint32_t j = 0; j < nList; ++j)
            weights[(size_t)t * ldW + j] = 0.f;
    if (S == 0) {
        free(alpha), free(beta), free(gamma);
        return FLT_MAX;
    }
    alpha[initial] = 0.0;
    for (uint32_t t = 0; t < T; ++t) {
        const float* row = scores + (size_t)t * ld;
        for (uint32_t s = 0; s < S; ++s) {
            const double from = alpha[(size_t)t * S + s];
            if (isinf(from))
                continue;
            for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                /* Aligner.cc:228-230: arcScore = transition + emission (f32), potential = score + arcScore */
                const float arcScore = arcWeight[a] + row[arcEmission[a]];
                double*     to       = &alpha[(size_t)(t + 1) * S + arcTarget[a]];
                *to                  = orc_log_add(*to, from + (double)arcScore);
            }
        }
    }
    double total = INFINITY;
    for (uint32_t s = 0; s < S; ++s)
        if (!isinf(finalWeight[s]))
            total = orc_log_add(total, alpha[(size_t)T * S + s] + (double)finalWeight[s]);
    if (isinf(total)) {
        free(alpha), free(beta), free(gamma);
        return FLT_MAX;
    }
    double *bNext = beta, *bCur = beta + S;
    for (uint32_t s = 0; s < S; ++s)
        bNext[s] = isinf(finalWeight[s]) ? INFINITY : (double)finalWeight[s];
    for (uint32_t t = T; t-- > 0;) {
        const float* row = scores + (size_t)t * ld;
        for (uint32_t j = 0; j < nList; ++j)
            gamma[j] = INFINITY;
        for (uint32_t s = 0; s < S; ++s) {
            bCur[s]           = INFINITY;
            const double from = alpha[(size_t)t * S + s];
            for (uint32_t a = arcOffsets[s]; a < arcOffsets[s + 1]; ++a) {
                const float  arcScore = arcWeight[a] + row[arcEmission[a]];
                const double rest     = (double)arcScore + bNext[arcTarget[a]];
                bCur[s]               = orc_log_add(bCur[s], rest);
                if (isinf(from) || isinf(rest))
                    continue;
                uint32_t j = 0;
                while (j < nList && emList[j] != arcEmission[a])
                    ++j;
                if (j < nList)
                    gamma[j] = orc_log_add(gamma[j], from + rest - total); /* combineItems, log semiring */
            }
        }
        double sum = 0.0; /* clip to >= 0, exp(-.), normalise over the frame (the shift cancels) */
        for (uint32_t j = 0; j < nList; ++j)
            sum += isinf(gamma[j]) ? 0.0 : exp(-(gamma[j] < 0.0 ? 0.0 : gamma[j]));
        for (uint32_t j = 0; j < nList; ++j)
            weights[(size_t)t * ldW + j] =
                    (isinf(gamma[j]) || sum == 0.0) ? 0.f : (float)(exp(-(gamma[j] < 0.0 ? 0.0 : gamma[j])) / sum);
        double* sw = bNext;
        bNext      = bCur;
        bCur       = sw;
    }
    free(alpha), free(beta), free(gamma);
    return total;
}
