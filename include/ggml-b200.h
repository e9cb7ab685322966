/*
 * ggml-b200.h — public C header of the B200 ggml backend plugin (libggml-b200.so).
 *
 * Takes the seat of ggml/include/ggml-hexagon.h:31-50 (ggml_backend_hexagon_init / _is_hexagon /
 * _get_device_count / _reg / _get_devname): same five entry points under a `b200` name, same meaning.
 * The plugin is loaded like any dynamic ggml backend: `ggml_backend_load("libggml-b200.so")`,
 * GGML_BACKEND_PATH, or discovery of libggml-*.so beside the executable (ggml-backend-reg.cpp:479-563);
 * the loader binds `ggml_backend_init` / `ggml_backend_score` (ggml-backend-impl.h:215-246).
 * Compile against the reference's own ggml.h / ggml-backend.h.
 */
#ifndef GGML_B200_H
#define GGML_B200_H

#include "ggml.h"
#include "ggml-backend.h"

#ifdef __cplusplus
extern "C" {
#endif

#define GGML_B200_MAX_DEVICES   8          /* one 8xB200 box (GGML_HEXAGON_MAX_DEVICES is 3) */
#define GGML_B200_BACKEND_NAME  "b200"

/* dev_num: CUDA ordinal. params: unused (ggml_backend_hexagon_init takes the QNN library path there). NULL on failure. */
GGML_BACKEND_API ggml_backend_t     ggml_backend_b200_init(size_t dev_num, const char * params);
GGML_BACKEND_API bool               ggml_backend_is_b200(ggml_backend_t backend);
GGML_BACKEND_API int                ggml_backend_b200_get_device_count(void);
GGML_BACKEND_API ggml_backend_reg_t ggml_backend_b200_reg(void);
GGML_BACKEND_API const char *       ggml_backend_b200_get_devname(size_t dev_num);

#ifdef __cplusplus
}
#endif
#endif /* GGML_B200_H */
