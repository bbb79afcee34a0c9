/* Deciding predecessor of every vertex in a MEASURED run of the tile flow (tools/flow_trace/analyse.py):
 * OUT -> the selected coverer that was popped first; IN -> the coverer popped last. Test/experiment tooling. */
#include <stdint.h>
void realpred(int64_t n, const int32_t* cp, const int32_t* ri, const double* len, double radius, const int64_t* pop,
              const uint8_t* in, int32_t* pred) {
    for (int64_t v = 0; v < n; ++v) {
        int64_t best_in = INT64_MAX, best_all = -1; int32_t pin = -1, pall = -1;
        for (int32_t a = cp[v]; a < cp[v+1]; ++a) {
            int32_t j = ri[a]; double d = len[a];
            if (!(d < radius)) continue;
            for (int32_t b = cp[j] - 1; b < cp[j+1]; ++b) {
                int32_t k; double dd;
                if (b < cp[j]) { k = j; dd = d; } else { k = ri[b]; dd = d + len[b]; }
                if (!(k < v && dd < radius)) continue;
                if (in[k] && pop[k] < best_in) { best_in = pop[k]; pin = k; }
                if (pop[k] > best_all) { best_all = pop[k]; pall = k; }
            }
        }
        pred[v] = in[v] ? pall : pin;
    }
}
