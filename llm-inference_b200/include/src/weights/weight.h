#pragma once
#include <string>
// reference src/weights/weight.h: abstract loader interface
struct Weight {
    virtual void loadWeights(std::string weight_path) = 0;
    virtual ~Weight() = default;
};
