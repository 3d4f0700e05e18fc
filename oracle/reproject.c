/*
 * oracle/reproject.c -- TEST INFRASTRUCTURE ONLY (see oracle/soft_ssv.c header).
 *
 * CPU restatement of the score-table contract: p-value -> per-model scale ->
 * int8 emission scores with the hit threshold at exactly 256. Parity: PINNED
 * against the reference's own PhmmReprojection.cpp compiled into
 * oracle/_ref/ (tests/test_oracle_vs_ref.py) and tests/golden/.
 * Compile with -ffp-contract=off: a fused multiply-add in the last step can
 * move a round() across a .5 boundary.
 */
#include <math.h>
#include <stdint.h>

/* PhmmReprojection/PhmmReprojection.cpp:15-31 (inverse Gumbel survival). */
double oracle_gumbel_invsurv(double p, double mu, double lambda)
{
    double lp = (p < 5e-9) ? (pow(p, p) - 1) / p : log(-1. * log(1 - p));
    return mu - (lp / lambda);
}

/* PhmmReprojection/PhmmReprojection.cpp:36-66. The float/double mix is part
 * of the contract: mu, lambda, maxLength, modelLength are floats; the Gumbel
 * score is a double; the N/C/B/E terms use logf; the null-model terms use the
 * C++ overload set: log(float) binds to the float overload (== logf) because
 * the reference is compiled as C++ and includes <math.h>, while
 * log(1.0 - float) has a double argument and binds to the double log. */
float oracle_scale_factor(float mu, float lambda, uint32_t max_length_u, uint32_t model_length_u,
                          float p_value)
{
    const float maxLength = (float)max_length_u;
    const float modelLength = (float)model_length_u;
    double fullModelScore = oracle_gumbel_invsurv(p_value, mu, lambda);
    float nLoop = logf(maxLength / (maxLength + 3));
    float nLoopTotal = nLoop * maxLength;
    float nEscape = logf(3.0f / (maxLength + 3));
    float bToM = logf(2.0f / (modelLength * (modelLength + 1)));
    float eToC = logf(1.0f / 2);
    float coreAdjust = (nEscape + nLoopTotal + nEscape + bToM + eToC);

    float bgLoop = maxLength / (maxLength + 1);
    float bgLoopTotal = maxLength * logf(bgLoop); /* C++: log(float) -> float overload */
    float bgMove = log(1.0 - bgLoop);
    float bgScore = bgLoopTotal + bgMove;

    float nats = (fullModelScore * 0.69314718055994529) + bgScore - coreAdjust;
    float bits = nats / 0.69314718055994529;
    return 256.0f / bits;
}

/* PhmmReprojection/PhmmReprojection.cpp:109-145: round(2m - s*(log2e*m)),
 * saturated to int8, for the 4*L negative-natural-log match emissions. */
void oracle_project_model(const float *match_emissions, uint32_t model_length, float scale,
                          int8_t *out)
{
    const float LOG2_E = 1.44269504089;
    const float alpha = 2 * scale;
    const float beta = LOG2_E * scale;
    float lengthAsFloat = (float)model_length;
    for (uint64_t i = 0; i < 4 * lengthAsFloat; i++) {
        float v = alpha - (match_emissions[i] * beta);
        v = round(v);
        if (v < -128) v = -128;
        if (v > 127) v = 127;
        out[i] = (int8_t)v;
    }
}
