#include <glm/gtx/rotate_vector.hpp>
