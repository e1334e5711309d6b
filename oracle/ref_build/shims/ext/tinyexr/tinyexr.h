// Stub for the tinyexr submodule, which is absent from the reference archive (test infrastructure
// only). No benchmark config loads an EXR image; every loader reports failure.
#pragma once

#define TINYEXR_SUCCESS 0
#define TINYEXR_ERROR_CANT_OPEN_FILE (-7)

static inline int LoadEXR(float **out_rgba, int *width, int *height, const char *, const char **err) {
    *out_rgba = nullptr;
    *width = 0;
    *height = 0;
    if (err) {
        *err = nullptr;
    }
    return TINYEXR_ERROR_CANT_OPEN_FILE;
}

static inline void FreeEXRErrorMessage(const char *) {}
