// Mirror of the reference's sift_cuda/types/CudaSiftConfig.hh:3-14 -- same struct, same (global) namespace,
// same field names (including the reference's spellings) and defaults, so caller code compiles unchanged.
#pragma once

struct CudaSiftConfig {
    int col_width;
    int row_width;
    int numFeatures {5000};
    int numOctaveLayers {3};             // only 3 is supported (as in the reference's kernels)
    double contrastThreshould {0.04};
    double edgeThreshould = 10;
    double sigma = 1.6;

    // The reference documents upscale=true as broken (CudaSiftConfig.hh:12); this engine rejects it.
    bool upscale = false;
};
