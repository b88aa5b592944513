/* window.c -- window generators, evaluated on the HOST at set-time exactly like the reference (signal/window.c:13-54):
 * the device only ever sees the finished table.  Generalised-cosine family: w[i] = alpha - (1-alpha)*cos(2*pi*i/D) with
 * the cosine term in double and the result narrowed to float; D = size-1 (symmetric) or size (periodic). */
#include <math.h>
#include "nntoolkitcore/signal/window.h"

static void cosine_window(float *w, int n, int period, float alpha) {
    const float beta = 1 - alpha; /* float subtraction, as in the reference */
    for (int i = 0; i < n; ++i) w[i] = (float)((double)alpha - (double)beta * cos(2 * M_PI * i / period));
}

void ones(float *vector, int size) { for (int i = 0; i < size; ++i) vector[i] = 1.0f; }
void hann_window(float *vector, int size) { cosine_window(vector, size, size - 1, 0.5f); }
void hamming_window(float *vector, int size) { cosine_window(vector, size, size - 1, 0.54f); }
void periodic_hann_window(float *vector, int size) { cosine_window(vector, size, size, 0.5f); }
void periodic_hamming_window(float *vector, int size) { cosine_window(vector, size, size, 0.54f); }
void blackman_window(float *vector, int size) {
    /* window.c:49-54: the phase is narrowed to float before the two double cosines */
    for (int i = 0; i < size; ++i) {
        float phase = (float)(2.f * M_PI * i / size);
        vector[i] = (float)(.42f - .5 * cos(phase) + .08 * cos(2 * phase));
    }
}
