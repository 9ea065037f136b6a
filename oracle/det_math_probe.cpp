// TEST INFRASTRUCTURE (oracle).  Exposes the project's deterministic
// transcendentals (mp_math.h) and libm's next to each other so tests can
// bound their difference.
#include <math.h>
#include "../madrona-mp-env_b200/csrc/mp_math.h"
extern "C" {
float probe_mp_sinf(float x) { return mpenv::mp_sinf(x); }
float probe_mp_cosf(float x) { return mpenv::mp_cosf(x); }
float probe_mp_logf(float x) { return mpenv::mp_logf(x); }
float probe_mp_atan2f(float y, float x) { return mpenv::mp_atan2f(y, x); }
float probe_mp_asinf(float x) { return mpenv::mp_asinf(x); }
float probe_libm_sinf(float x) { return sinf(x); }
float probe_libm_cosf(float x) { return cosf(x); }
float probe_libm_logf(float x) { return logf(x); }
float probe_libm_atan2f(float y, float x) { return atan2f(y, x); }
float probe_libm_asinf(float x) { return asinf(x); }
void probe_batch(int which, int n, const float *a, const float *b, float *mp, float *lm)
{
    for (int i = 0; i < n; i++) {
        switch (which) {
        case 0: mp[i] = mpenv::mp_sinf(a[i]); lm[i] = sinf(a[i]); break;
        case 1: mp[i] = mpenv::mp_cosf(a[i]); lm[i] = cosf(a[i]); break;
        case 2: mp[i] = mpenv::mp_logf(a[i]); lm[i] = logf(a[i]); break;
        case 3: mp[i] = mpenv::mp_atan2f(a[i], b[i]); lm[i] = atan2f(a[i], b[i]); break;
        case 4: mp[i] = mpenv::mp_asinf(a[i]); lm[i] = asinf(a[i]); break;
        }
    }
}
}
