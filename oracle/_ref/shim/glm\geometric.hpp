#include <glm/geometric.hpp>
