/* Test-infrastructure shim (oracle/): nori/common.h:39 includes <ImathPlatform.h> from
 * IlmBase only for platform macros; nothing on the render path uses it. IlmBase is not
 * installed in this image, so the oracle build supplies this empty stand-in. */
#pragma once
