/* TEST INFRASTRUCTURE.  Empty stand-in for visionaray's private <common/image.h>, which the reference's
 * src/vkt/Render_kernel.hpp:13 includes only for code that is commented out (:69-76) and which would pull in Boost
 * (absent from this image).  The reference's kernel source itself is compiled unmodified. */
