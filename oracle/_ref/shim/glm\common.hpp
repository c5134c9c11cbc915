#include <glm/common.hpp>
