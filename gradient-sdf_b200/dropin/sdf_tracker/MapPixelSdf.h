// Drop-in for cpp/include/sdf_tracker/MapPixelSdf.h (:52-155): the plain SDF volume, constructors :97-101.
#ifndef GSDF_DROPIN_MAP_PIXEL_SDF_H_
#define GSDF_DROPIN_MAP_PIXEL_SDF_H_
#include "Sdf.h"
class MapPixelSdf : public Sdf {
public:
    MapPixelSdf() : Sdf(new gsdf_shell::MapPixelSdf(), 0.02f) {}
    explicit MapPixelSdf(float voxel_size) : Sdf(new gsdf_shell::MapPixelSdf(voxel_size), voxel_size) {}
    MapPixelSdf(float voxel_size, float T) : Sdf(new gsdf_shell::MapPixelSdf(voxel_size, T), voxel_size) {}
};
#endif
