// engine_internal.h -- what other translation units of libfjpmml.so may ask the engine about its opaque handles.
#pragma once
#include "../../include/fjpmml.h"
#include "model.h"

namespace fjp {
const DeviceModel* internal_device_model(const fjp_model* m);   // NULL when the model is not device-resident
int internal_model_device(const fjp_model* m);
int internal_model_fields(const fjp_model* m);
int internal_model_derived(const fjp_model* m);                  // number of DerivedFields (they need extra scratch)
void internal_count_launches(fjp_engine* e, int launches);
void internal_set_error(const char* msg);                        // text returned by fjp_last_error() on this thread
}  // namespace fjp
