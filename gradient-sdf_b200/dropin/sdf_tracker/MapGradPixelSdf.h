// Drop-in for cpp/include/sdf_tracker/MapGradPixelSdf.h (:51-152): the Gradient-SDF volume, constructors :99-103.
#ifndef GSDF_DROPIN_MAP_GRAD_PIXEL_SDF_H_
#define GSDF_DROPIN_MAP_GRAD_PIXEL_SDF_H_
#include "Sdf.h"
class MapGradPixelSdf : public Sdf {
public:
    MapGradPixelSdf() : Sdf(new gsdf_shell::MapGradPixelSdf(), 0.02f) {}
    explicit MapGradPixelSdf(float voxel_size) : Sdf(new gsdf_shell::MapGradPixelSdf(voxel_size), voxel_size) {}
    MapGradPixelSdf(float voxel_size, float T) : Sdf(new gsdf_shell::MapGradPixelSdf(voxel_size, T), voxel_size) {}
};
#endif
