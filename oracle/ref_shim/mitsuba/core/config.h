// TEST INFRASTRUCTURE: build configuration for compiling the reference in place (oracle/Makefile.ref).
// Written for this repo: it plays the role of the header the reference build derives from mitsuba.conf
// (variants enabled here: scalar_rgb -- needed to link libmitsuba-core --, scalar_mono, scalar_mono_double).
#pragma once
#include <mitsuba/core/fwd.h>

#define MTS_VARIANTS                                                            \
    "scalar_rgb\n"                                                              \
    "scalar_mono\n"                                                             \
    "scalar_mono_double\n"                                                      \

#define MTS_DEFAULT_VARIANT "scalar_mono"

#define MTS_EXTERN_STRUCT_CORE(Name)                                            \
    MTS_EXTERN_CORE template struct Name<float, Color<float, 3>>;               \
    MTS_EXTERN_CORE template struct Name<float, Color<float, 1>>;               \
    MTS_EXTERN_CORE template struct Name<double, Color<double, 1>>;             \

#define MTS_EXTERN_CLASS_CORE(Name)                                             \
    MTS_EXTERN_CORE template class Name<float, Color<float, 3>>;                \
    MTS_EXTERN_CORE template class Name<float, Color<float, 1>>;                \
    MTS_EXTERN_CORE template class Name<double, Color<double, 1>>;              \

#define MTS_EXTERN_STRUCT_RENDER(Name)                                          \
    MTS_EXTERN_RENDER template struct Name<float, Color<float, 3>>;             \
    MTS_EXTERN_RENDER template struct Name<float, Color<float, 1>>;             \
    MTS_EXTERN_RENDER template struct Name<double, Color<double, 1>>;           \

#define MTS_EXTERN_CLASS_RENDER(Name)                                           \
    MTS_EXTERN_RENDER template class Name<float, Color<float, 3>>;              \
    MTS_EXTERN_RENDER template class Name<float, Color<float, 1>>;              \
    MTS_EXTERN_RENDER template class Name<double, Color<double, 1>>;            \

#define MTS_INSTANTIATE_STRUCT(Name)                                            \
    template struct Name<float, Color<float, 3>>;                               \
    template struct Name<float, Color<float, 1>>;                               \
    template struct Name<double, Color<double, 1>>;                             \

#define MTS_INSTANTIATE_CLASS(Name)                                             \
    template class Name<float, Color<float, 3>>;                                \
    template class Name<float, Color<float, 1>>;                                \
    template class Name<double, Color<double, 1>>;                              \

#define MTS_INVOKE_VARIANT(variant, func, ...)                                  \
    [&]() {                                                                     \
        if (variant == "scalar_rgb")                                            \
            return func<float, Color<float, 3>>(__VA_ARGS__);                   \
        else if (variant == "scalar_mono")                                      \
            return func<float, Color<float, 1>>(__VA_ARGS__);                   \
        else if (variant == "scalar_mono_double")                               \
            return func<double, Color<double, 1>>(__VA_ARGS__);                 \
        else                                                                    \
            Throw("Unsupported variant: %s", variant);                          \
    }()                                                                         \

NAMESPACE_BEGIN(mitsuba)
NAMESPACE_BEGIN(detail)
template <typename Float_, typename Spectrum_> constexpr const char *get_variant() {
    if constexpr (std::is_same_v<Float_, float> && std::is_same_v<Spectrum_, Color<float, 3>>)
        return "scalar_rgb";
    else if constexpr (std::is_same_v<Float_, float> && std::is_same_v<Spectrum_, Color<float, 1>>)
        return "scalar_mono";
    else if constexpr (std::is_same_v<Float_, double> && std::is_same_v<Spectrum_, Color<double, 1>>)
        return "scalar_mono_double";
    else
        return "";
}
NAMESPACE_END(detail)
NAMESPACE_END(mitsuba)
