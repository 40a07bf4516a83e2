// handle.hpp -- what a tk_state handle holds (shared by capi.cu and backend_local.cu).
#pragma once
#include <string>
#include <string_view>
#include <utility>
#include <vector>

#include "doc.hpp"
#include "engine.hpp"

struct tk_state {
    std::string name;
    tk::Document doc;
    tk::Engine eng;
    std::string err;
    std::string blob;  // last Clusters()/Nodes() result
    std::string prefix;
    std::vector<std::pair<std::string_view, std::string_view>> pairs;
};
