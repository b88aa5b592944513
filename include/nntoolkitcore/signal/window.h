/*
 * nntoolkitcore/signal/window.h -- window generators (B200 build of NNToolkitCore's inference path).
 *
 * Drop-in for the reference header signal/window.h:17-29.  Windows are evaluated ON THE HOST at set time, in
 * double precision rounded once to float, exactly as the reference does (signal/window.c:13-17,49-54); the
 * device only ever sees the finished table (it is uploaded by SpectrogramSetWindowFunc).  A user-supplied
 * window_fn therefore keeps working unchanged.
 */
#ifndef NNTK_B200_SIGNAL_WINDOW_H
#define NNTK_B200_SIGNAL_WINDOW_H

#ifdef __cplusplus
extern "C" {
#endif

/* Fills vector[0..size) with the window.  Reference: signal/window.h:17. */
typedef void (*window_fn)(float *vector, int size);

void ones(float *vector, int size);                     /* all 1.0f                    -- window.c:27-31 */
void hann_window(float *vector, int size);              /* symmetric, D = size-1, 0.5  -- window.c:45-47 */
void hamming_window(float *vector, int size);           /* symmetric, D = size-1, 0.54 -- window.c:41-43 */
void periodic_hann_window(float *vector, int size);     /* periodic,  D = size,   0.5  -- window.c:37-39 */
void periodic_hamming_window(float *vector, int size);  /* periodic,  D = size,   0.54 -- window.c:33-35 */
void blackman_window(float *vector, int size);          /* periodic 3-term Blackman    -- window.c:49-54 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SIGNAL_WINDOW_H */
