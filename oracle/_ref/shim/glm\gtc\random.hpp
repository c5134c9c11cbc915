#include <glm/gtc/random.hpp>
