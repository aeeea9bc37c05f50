package ai.langstream.cuda;

import java.io.IOException;
import java.io.InputStream;
import java.nio.file.Files;
import java.nio.file.Path;
import java.nio.file.StandardCopyOption;

/**
 * NAR archives are unpacked and loaded through a URLClassLoader (NarFileHandler.java:123-130,174-183),
 * and a JNI library can be bound to one classloader only. Like AbstractHuggingFaceEmbeddingService's
 * static initialiser for libtorch (:45-49), extract the libraries once per JVM and System.load them.
 */
final class NativeLoader {
    private static boolean loaded;

    private NativeLoader() {}

    static synchronized void loadOnce(String... libs) {
        if (loaded) {
            return;
        }
        try {
            Path dir = Files.createTempDirectory("langstream-cuda");
            for (String lib : libs) {
                String file = "lib" + lib + ".so";
                try (InputStream in = NativeLoader.class.getResourceAsStream("/native/" + file)) {
                    if (in == null) {
                        throw new IllegalStateException(file + " is not packaged in this NAR");
                    }
                    Path target = dir.resolve(file);
                    Files.copy(in, target, StandardCopyOption.REPLACE_EXISTING);
                    System.load(target.toAbsolutePath().toString());
                }
            }
            loaded = true;
        } catch (IOException e) {
            throw new IllegalStateException("cannot extract the langstream-cuda native libraries", e);
        } catch (UnsatisfiedLinkError e) {
            throw new IllegalStateException(
                    "langstream-cuda was already loaded by another classloader, or no usable GPU driver", e);
        }
    }
}
